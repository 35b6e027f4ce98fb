/* Test infrastructure only.  press.c references ~30 third-party entry points
 * (bzip2, fast-lzma2, FLAC, TurboPFor, Turbo-Range-Coder, sigtk/jnn) that are
 * all OFF the hot path.  To link the
 * reference's own press.c unmodified into oracle/_ref/libpress_ref.so we
 * define them here as traps: calling one is a harness bug. */
#include <stdio.h>
#include <stdlib.h>
#define TRAP(name) void name(void) { fprintf(stderr, "oracle/_ref: off-path symbol " #name " called\n"); abort(); }
TRAP(BZ2_bzBuffToBuffCompress) TRAP(BZ2_bzBuffToBuffDecompress)
TRAP(FL2_compressBound) TRAP(FL2_compressMt) TRAP(FL2_decompressMt) TRAP(FL2_isError)
TRAP(FLAC__stream_decoder_delete) TRAP(FLAC__stream_decoder_init_stream)
TRAP(FLAC__stream_decoder_new) TRAP(FLAC__stream_decoder_process_until_end_of_stream)
TRAP(FLAC__stream_encoder_delete) TRAP(FLAC__stream_encoder_finish)
TRAP(FLAC__stream_encoder_init_stream) TRAP(FLAC__stream_encoder_new)
TRAP(FLAC__stream_encoder_process_interleaved) TRAP(FLAC__stream_encoder_set_bits_per_sample)
TRAP(FLAC__stream_encoder_set_channels) TRAP(FLAC__stream_encoder_set_compression_level)
TRAP(FLAC__stream_encoder_set_do_exhaustive_model_search) TRAP(FLAC__stream_encoder_set_sample_rate)
TRAP(FLAC__stream_encoder_set_total_samples_estimate)
TRAP(jnn_raw) TRAP(p4nzdec128v16) TRAP(p4nzenc128v16)
TRAP(rccdfdec) TRAP(rccdfenc) TRAP(rccsdec) TRAP(rccsenc)
TRAP(rcmsdec) TRAP(rcmsenc) TRAP(rcsdec) TRAP(rcsenc)
