/* API version this library is a drop-in for (reference include/defl/version.h:20-22). */
#ifndef defl_version_h
#define defl_version_h

#define DEFL_VERSION_MAJOR 0
#define DEFL_VERSION_MINOR 1
#define DEFL_VERSION_PATCH 1

/* version of the B200 engine behind that API */
#define DEFL_B200_VERSION_MAJOR 0
#define DEFL_B200_VERSION_MINOR 1

#endif
