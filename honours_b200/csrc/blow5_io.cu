// blow5_io.cu -- the steps either side of the hot path: BLOW5 records in,
// an on-disk container out.  Host code only (compiled with the library; nothing here is a kernel).
//
//   reader     replaces, for the reference's benchmark loop, slow5_open / slow5_get_next
//              (press/test.h:10-37; press/sigtk/slow5lib/src/slow5.c:805-830 binary header, :3314
//              slow5_get_next, :3925-3960 record layout).  Records are inflated on the host (zlib) and
//              their signals land in the batch layout of include/press_gpu.h (`samples`, `offsets`).
//              When the file keeps its signals svb-zd coded (signal-press byte 1: the field is
//              [n u32][classical streamvbyte over zd], slow5_press.c), the stored frames are NOT
//              decoded on the host: they are exactly PRESS_GPU_VBZ_SVB_ZD streams and go through
//              press_gpu_batch_depress_host, i.e. the GPU.
//   container  [magic][method][nreads][nsamples][nbytes][sample offsets][stream offsets][payload]:
//              what press_gpu_batch_press_host returns, on disk, and back.
#include "../../include/press_gpu.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#include <string>
#include <vector>

namespace {

const char BLOW5_MAGIC[6] = {'B', 'L', 'O', 'W', '5', '\1'};
const char BLOW5_EOF[5] = {'5', 'W', 'O', 'L', 'B'};
const char CONTAINER_MAGIC[8] = {'P', 'R', 'S', 'G', 'P', 'U', '1', '\0'};

void io_err(const char *fmt, const char *what)
{
	fprintf(stderr, "press_gpu: ");
	fprintf(stderr, fmt, what);
	fprintf(stderr, "\n");
}

}  // namespace

struct press_gpu_blow5 {
	FILE *fp = nullptr;
	uint8_t record_press = 0;	// 0 none, 1 zlib (slow5_press.h)
	uint8_t signal_press = 0;	// 0 none, 1 svb-zd
	bool eof = false;
	std::vector<uint8_t> zbuf, rec;
	// a record read ahead that did not fit the previous batch
	bool pending = false;
	std::string pending_id;
	std::vector<uint8_t> pending_sig;	// raw samples or the svb-zd frame
	uint64_t pending_n = 0;
};

namespace {

// one record -> (id, signal field, sample count); false at the end-of-file marker or on error
bool next_record(press_gpu_blow5 *f, std::string &id, std::vector<uint8_t> &sig, uint64_t &n)
{
	if (f->eof)
		return false;
	uint8_t head[8];
	if (fread(head, 1, 5, f->fp) != 5) {
		f->eof = true;
		return false;
	}
	if (!memcmp(head, BLOW5_EOF, 5)) {
		f->eof = true;
		return false;
	}
	if (fread(head + 5, 1, 3, f->fp) != 3) {
		f->eof = true;
		return false;
	}
	uint64_t size;
	memcpy(&size, head, 8);
	f->zbuf.resize(size);
	if (fread(f->zbuf.data(), 1, size, f->fp) != size) {
		io_err("blow5: truncated record%s", "");
		f->eof = true;
		return false;
	}
	const uint8_t *rec = f->zbuf.data();
	uint64_t rlen = size;
	if (f->record_press == 1) {	// zlib, slow5_press.c
		size_t cap = f->rec.size() < 4 * size + 4096 ? 4 * size + 4096 : f->rec.size();
		for (;;) {
			f->rec.resize(cap);
			uLongf dl = (uLongf) cap;
			const int z = uncompress(f->rec.data(), &dl, f->zbuf.data(), (uLong) size);
			if (z == Z_OK) {
				rlen = dl;
				break;
			}
			if (z != Z_BUF_ERROR) {
				io_err("blow5: zlib failed on a record%s", "");
				f->eof = true;
				return false;
			}
			cap *= 2;
		}
		rec = f->rec.data();
	} else if (f->record_press != 0) {
		io_err("blow5: record compression method not supported (only none / zlib)%s", "");
		f->eof = true;
		return false;
	}
	// [id length u16][id][read group u32][digitisation, offset, range, sampling rate f64 x 4][length u64][signal]
	uint16_t idlen;
	if (rlen < 2)
		return false;
	memcpy(&idlen, rec, 2);
	const uint64_t p = 2 + (uint64_t) idlen + 4 + 4 * 8;
	if (rlen < p + 8)
		return false;
	id.assign(reinterpret_cast<const char *>(rec + 2), idlen);
	uint64_t field;
	memcpy(&field, rec + p, 8);
	const uint64_t nbytes = f->signal_press ? field : 2 * field;	// (svb-zd: the field is the frame's byte length)
	if (rlen < p + 8 + nbytes)
		return false;
	sig.assign(rec + p + 8, rec + p + 8 + nbytes);
	if (f->signal_press) {
		uint32_t cnt = 0;
		if (nbytes >= 4)
			memcpy(&cnt, sig.data(), 4);
		n = cnt;
	} else {
		n = field;
	}
	return true;
}

}  // namespace

extern "C" {

press_gpu_blow5 *press_gpu_blow5_open(const char *path)
{
	FILE *fp = fopen(path, "rb");
	if (!fp) {
		io_err("blow5: cannot open %s", path);
		return nullptr;
	}
	uint8_t h[68];
	if (fread(h, 1, 68, fp) != 68 || memcmp(h, BLOW5_MAGIC, 6)) {
		io_err("blow5: %s is not a BLOW5 file", path);
		fclose(fp);
		return nullptr;
	}
	press_gpu_blow5 *f = new press_gpu_blow5;
	f->fp = fp;
	f->record_press = h[9];
	// the signal-press byte exists from file version 0.2.0 on (slow5.c:820)
	const bool has_sig = h[6] > 0 || h[7] >= 2;
	f->signal_press = has_sig ? h[14] : 0;
	uint32_t hdr_len;
	memcpy(&hdr_len, h + 64, 4);
	if (f->signal_press > 1 || fseek(fp, 68 + (long) hdr_len, SEEK_SET) != 0) {
		io_err("blow5: %s: unsupported signal compression or short header", path);
		fclose(fp);
		delete f;
		return nullptr;
	}
	return f;
}

void press_gpu_blow5_close(press_gpu_blow5 *f)
{
	if (!f)
		return;
	if (f->fp)
		fclose(f->fp);
	delete f;
}

int press_gpu_blow5_signal_press(const press_gpu_blow5 *f) { return f ? f->signal_press : -1; }

int64_t press_gpu_blow5_next_batch(press_gpu_blow5 *f, uint64_t max_reads, uint64_t max_samples, int16_t *h_samples,
				   uint64_t *h_offsets, char *ids, size_t id_stride)
{
	if (!f || !h_samples || !h_offsets)
		return -1;
	std::vector<uint8_t> frames;		// svb-zd files: the stored frames, back to back
	std::vector<uint64_t> frame_off(1, 0);
	uint64_t nreads = 0, ns = 0;
	h_offsets[0] = 0;
	while (nreads < max_reads) {
		std::string id;
		std::vector<uint8_t> sig;
		uint64_t n = 0;
		if (f->pending) {
			id.swap(f->pending_id);
			sig.swap(f->pending_sig);
			n = f->pending_n;
			f->pending = false;
		} else if (!next_record(f, id, sig, n)) {
			break;
		}
		if (ns + n > max_samples) {	// keep it for the next batch
			if (nreads == 0) {
				io_err("blow5: a read does not fit the batch buffer%s", "");
				return -1;
			}
			f->pending = true;
			f->pending_id.swap(id);
			f->pending_sig.swap(sig);
			f->pending_n = n;
			break;
		}
		if (f->signal_press) {
			frames.insert(frames.end(), sig.begin(), sig.end());
			frame_off.push_back(frames.size());
		} else {
			memcpy(h_samples + ns, sig.data(), 2 * n);
		}
		if (ids && id_stride) {
			const size_t c = id.size() < id_stride - 1 ? id.size() : id_stride - 1;
			memcpy(ids + nreads * id_stride, id.data(), c);
			ids[nreads * id_stride + c] = '\0';
		}
		ns += n;
		h_offsets[++nreads] = ns;
	}
	if (nreads && f->signal_press) {
		// the stored frames are pre-zstd VBZ streams ([n u32][svb over zd]): decoded by the GPU
		frames.resize(frames.size() + 64);
		if (press_gpu_batch_depress_host(PRESS_GPU_VBZ_SVB_ZD, frames.data(), frame_off.data(), h_offsets, nreads,
						 h_samples) != 0)
			return -1;
	}
	return (int64_t) nreads;
}

/* ---- container ------------------------------------------------------------------------------- */

int press_gpu_container_write(const char *path, int method, uint64_t nreads, const uint64_t *h_offsets,
			      const uint64_t *h_out_offsets, const uint8_t *h_container)
{
	FILE *fp = fopen(path, "wb");
	if (!fp) {
		io_err("container: cannot create %s", path);
		return -1;
	}
	const uint64_t hdr[4] = {(uint64_t) (uint32_t) method, nreads, h_offsets[nreads] - h_offsets[0],
				 h_out_offsets[nreads] - h_out_offsets[0]};
	bool ok = fwrite(CONTAINER_MAGIC, 1, 8, fp) == 8 && fwrite(hdr, 8, 4, fp) == 4;
	for (uint64_t r = 0; ok && r <= nreads; r++) {
		const uint64_t v = h_offsets[r] - h_offsets[0];
		ok = fwrite(&v, 8, 1, fp) == 1;
	}
	for (uint64_t r = 0; ok && r <= nreads; r++) {
		const uint64_t v = h_out_offsets[r] - h_out_offsets[0];
		ok = fwrite(&v, 8, 1, fp) == 1;
	}
	ok = ok && fwrite(h_container + h_out_offsets[0], 1, hdr[3], fp) == hdr[3];
	ok = fclose(fp) == 0 && ok;
	if (!ok)
		io_err("container: write to %s failed", path);
	return ok ? 0 : -1;
}

int press_gpu_container_info(const char *path, int *method, uint64_t *nreads, uint64_t *nsamples, uint64_t *nbytes)
{
	FILE *fp = fopen(path, "rb");
	char magic[8];
	uint64_t hdr[4];
	const bool ok = fp && fread(magic, 1, 8, fp) == 8 && !memcmp(magic, CONTAINER_MAGIC, 8) && fread(hdr, 8, 4, fp) == 4;
	if (fp)
		fclose(fp);
	if (!ok) {
		io_err("container: %s is not a press_gpu container", path);
		return -1;
	}
	*method = (int) hdr[0];
	*nreads = hdr[1];
	*nsamples = hdr[2];
	*nbytes = hdr[3];
	return 0;
}

int press_gpu_container_read(const char *path, uint64_t *h_offsets, uint64_t *h_out_offsets, uint8_t *h_container)
{
	int method;
	uint64_t nreads, nsamples, nbytes;
	if (press_gpu_container_info(path, &method, &nreads, &nsamples, &nbytes))
		return -1;
	FILE *fp = fopen(path, "rb");
	bool ok = fp && fseek(fp, 8 + 32, SEEK_SET) == 0 && fread(h_offsets, 8, nreads + 1, fp) == nreads + 1 &&
		  fread(h_out_offsets, 8, nreads + 1, fp) == nreads + 1 && fread(h_container, 1, nbytes, fp) == nbytes;
	if (fp)
		fclose(fp);
	ok = ok && h_offsets[nreads] == nsamples && h_out_offsets[nreads] == nbytes;
	for (uint64_t r = 0; ok && r < nreads; r++)
		ok = h_offsets[r] <= h_offsets[r + 1] && h_out_offsets[r] <= h_out_offsets[r + 1];
	if (!ok)
		io_err("container: %s is truncated or inconsistent", path);
	return ok ? 0 : -1;
}

}  // extern "C"
