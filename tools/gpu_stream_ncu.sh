#!/bin/bash
# one ncu --set full capture of the stream kernel (merge row phase) on RMAT-22
mkdir -p gpurun_out
export SHB_STREAM_MERGE=${SHB_STREAM_MERGE:-1}
timeout 200 python tools/stream_bench.py > gpurun_out/stream_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmv_stream -s 3 -c 1 -o gpurun_out/prof_stream_merge python tools/stream_bench.py > gpurun_out/ncu_stream.log 2>&1
echo "rc=$?"; tail -1 gpurun_out/stream_plain.log; ls -la gpurun_out | grep prof_stream
