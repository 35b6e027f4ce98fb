/* Declaration-only stand-in for <zstd.h> (the image ships libzstd.so.1 but no
 * header).  Only the five entry points press.c names are declared; they are
 * resolved against the real libzstd.so.1 at link time (see oracle/Makefile).
 * Test infrastructure only: used to compile the reference in place into
 * oracle/_ref/.  zstd itself is outside the hot path. */
#ifndef ORACLE_ZSTD_DECLS_H
#define ORACLE_ZSTD_DECLS_H
#include <stddef.h>
size_t ZSTD_compressBound(size_t srcSize);
size_t ZSTD_compress(void *dst, size_t dstCapacity, const void *src,
		     size_t srcSize, int compressionLevel);
size_t ZSTD_decompress(void *dst, size_t dstCapacity, const void *src,
		       size_t compressedSize);
unsigned ZSTD_isError(size_t code);
unsigned long long ZSTD_getFrameContentSize(const void *src, size_t srcSize);
#endif
