// common.h -- shared bits of the harness host (counterpart of the reference's inc/common.h:1-56).
//
// The reference packs the whole start of every main() into the COMMON_MAIN_PREAMBLE macro
// (inc/common.h:5-56).  Here the same steps (same flags, same defaults, same messages, same exit
// codes) live in one function template, harness_preamble<T>() in app_common.h, which returns
// a struct instead of injecting locals.
#pragma once

#define ENDL "\n"
