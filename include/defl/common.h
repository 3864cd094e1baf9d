/*
 * defl/common.h -- result codes, flags and handle types of the inflate API.
 *
 * Drop-in for the reference header include/defl/common.h: same names and the same
 * numeric values (they are ABI): UnzResult (common.h:80-90), INFL_ZLIB (common.h:99),
 * the opaque infl_stream_t (common.h:94-96) and the UNZ_EXPORT / UNZ_INLINE macros
 * (common.h:50-74).  The stream layout behind the handle is private to this library.
 */
#ifndef defl_common_h
#define defl_common_h

#include <stdint.h>
#include <stddef.h>
#include <stdbool.h>
#include <errno.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_MSC_VER) || defined(__MINGW32__) || defined(__MINGW64__)
#  if defined(UNZ_STATIC)
#    define UNZ_EXPORT
#  elif defined(UNZ_EXPORTS)
#    define UNZ_EXPORT __declspec(dllexport)
#  else
#    define UNZ_EXPORT __declspec(dllimport)
#  endif
#  define UNZ_HIDE
#  define UNZ_INLINE __forceinline
#else
#  define UNZ_EXPORT __attribute__((visibility("default")))
#  define UNZ_HIDE   __attribute__((visibility("hidden")))
#  define UNZ_INLINE static inline __attribute__((always_inline))
#endif

#ifndef unz_common_h
typedef enum UnzResult {
  UNZ_UNFINISHED =  2,        /* infl_stream(): feed more input              */
  UNZ_NOOP       =  1,        /* nothing to do (no chunk was ever included)  */
  UNZ_OK         =  0,
  UNZ_ERR        = -1,        /* malformed / truncated input, or dst too small in infl() */
  UNZ_EFOUND     = -1000,
  UNZ_ENOMEM     = -ENOMEM,
  UNZ_EPERM      = -EPERM,
  UNZ_EBADF      = -EBADF,
  UNZ_EFULL      = -ENOBUFS   /* infl_stream(): dst full, call infl_resize_output() */
} UnzResult;
#endif

typedef struct unz__stream_t defl_stream_t;
typedef struct unz__stream_t infl_stream_t;
typedef struct unz__chunk_t  defl_chunk_t;

/* inflate flags: compared by EQUALITY, as the reference does (src/infl/infl.c:224,1026,1565):
   1 = zlib wrapper, anything else = raw deflate */
#define INFL_ZLIB 1

#ifdef __cplusplus
}
#endif
#endif /* defl_common_h */
