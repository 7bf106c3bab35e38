#!/bin/bash
# quick GPU call: run the tests named on the command line
timeout 900 python -m pytest "$@" -m gpu -x -q --timeout 600 2>&1 | tail -15
