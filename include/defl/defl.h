/* The reference's compressor header is an empty placeholder (include/defl/defl.h:17-25,
   README.md:234); kept so `#include <defl/defl.h>` keeps compiling. */
#ifndef defl_h
#define defl_h
#include "common.h"
#endif
