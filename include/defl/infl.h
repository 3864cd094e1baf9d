/*
 * defl/infl.h -- the inflate API of recp/defl, served by the B200 engine.
 *
 * Every entry point below replaces the reference function of the same name and
 * signature (reference: include/defl/infl.h, line numbers cited per function).
 * Behavioural contract: infl()/infl_buf() return UNZ_OK, UNZ_ERR or
 * UNZ_NOOP only -- a too-small dst is UNZ_ERR, not UNZ_EFULL (src/infl/infl.c:1078-1087);
 * only infl_stream() returns UNZ_UNFINISHED / UNZ_EFULL.  The zlib trailer is not
 * required and not verified on this API (the reference never reads it); Adler-32 is
 * computed on the device and available through infl_adler32() and the batch API
 * (defl/infl_batch.h).
 *
 * There is no CPU decode path: every call needs a CUDA device and fails with UNZ_ERR
 * (infl_init: NULL) when none is usable.
 */
#ifndef infl_h
#define infl_h

#include "common.h"

#ifdef __cplusplus
extern "C" {
#endif

/* reference infl.h:31-33 -- new one-shot/streaming handle writing into dst[0..dstlen) */
UNZ_EXPORT infl_stream_t *infl_init(void *dst, uint32_t dstlen, int flags);

/* reference infl.h:43-48 -- reuse a handle for a new destination */
UNZ_EXPORT void infl_reset(infl_stream_t *stream, void *dst, uint32_t dstlen, int flags);

/* reference infl.h:60-64 -- swap the output buffer after UNZ_EFULL, keeping state */
UNZ_EXPORT int infl_resize_output(infl_stream_t *stream, void *dst, uint32_t dstlen);

/* reference infl.h:78-82 -- append one piece of compressed input; the logical input is the
   concatenation of all pieces in call order (src/infl/mem.c:114-216).  Pieces up to 8 KiB
   are copied (the caller may release them at once); larger pieces are borrowed until
   infl() returns (README.md:47). */
UNZ_EXPORT void infl_include(infl_stream_t *stream, const void *ptr, uint32_t len);

/* reference infl.h:92-94 -- decode everything included so far */
UNZ_EXPORT int infl(infl_stream_t *stream);

/* reference infl.h:102-104 -- drop the included pieces, keep the handle */
UNZ_EXPORT void infl_reset_pool(infl_stream_t *stream);

/* reference infl.h:111-113 -- NULL is a no-op */
UNZ_EXPORT void infl_destroy(infl_stream_t *stream);

/* reference infl.h:125-145 (header-inline there; exported here so FFI callers can bind it) */
UNZ_EXPORT int infl_buf(const void *src, uint32_t srclen, void *dst, uint32_t dstlen, int flags);

/* reference infl.h:164-168 -- resumable decode; see INTEGRATION.md for the host shim */
UNZ_EXPORT int infl_stream(infl_stream_t *stream, const void *src, uint32_t srclen);

/* reference infl.h:175-177 */
UNZ_EXPORT uint32_t infl_output_pos(const infl_stream_t *stream);

/* reference infl.h:187-189 -- compressed bytes consumed; stops right before a zlib trailer */
UNZ_EXPORT uint32_t infl_input_pos(const infl_stream_t *stream);

/* NEW (not in the reference): Adler-32 of the bytes produced by the last successful
   infl()/infl_stream() on this handle, and the big-endian trailer found after the body
   (0 and *present = 0 when the input ends before it). */
UNZ_EXPORT uint32_t infl_adler32(const infl_stream_t *stream, uint32_t *stored, int *present);

#ifdef __cplusplus
}
#endif
#endif /* infl_h */
