// fitsread.cu — raw FITS data units into a resident frame block.
//
// Replaces, for uncompressed primary-HDU images, the read leg of stack_read_block_data
// (median_and_mean.c:382-545): seq_opened_read_region -> internal_read_partial_fits
// (image_format_fits.c:927-1015), i.e. cfitsio's fits_read_subset (big-endian decode, BZERO) followed
// by convert_data_ushort (:390-416) / convert_floats (:648-672).  The host only moves bytes: file
// rows go pread -> pinned staging -> H2D, two chunks in flight, and the decode (byte swap, +32768,
// BYTE widening, the DATAMAX > 10 rescale, y shift, zero fill) is a kernel that writes straight into
// the block.  Rows of a FITS data unit are already in the order Siril keeps in memory (bottom-up),
// so frame row r of layer l is at byte ((l * frame_height + r) * width) * |BITPIX| / 8.
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <fcntl.h>
#include <unistd.h>
#include <thread>
#include <functional>
#include <sys/stat.h>
#include "common.cuh"

#define SGPU_STAGE_CHUNK ((size_t) 32 << 20)	// bytes of raw rows per staging buffer
#define INV_USHRT_MAX_SINGLE .000015259022f	// core/siril.h:94

// ---- device decode -------------------------------------------------------------------------------
// 16-bit: big-endian int16 + 32768 as WORD.  USHORT_IMG (BZERO 32768 applied by cfitsio) and
// SHORT_IMG (convert_data_ushort: 32768 + (int) data16[i]) are the same map: swap, flip the top bit.
__device__ __forceinline__ uint32_t dec16x2(uint32_t w) { return __byte_perm(w, 0, 0x2301) ^ 0x80008000u; }
__device__ __forceinline__ uint32_t bswap32(uint32_t w) { return __byte_perm(w, 0, 0x0123); }
__device__ __forceinline__ uint32_t dec32(uint32_t w, bool rescale) {
	w = bswap32(w);
	return rescale ? __float_as_uint(__uint_as_float(w) * INV_USHRT_MAX_SINGLE) : w;	// convert_floats, default case
}

// raw and dst have the same address modulo 16 (the staging offset is chosen so): scalar head up to the
// first 16-byte boundary, 16-byte body, scalar tail.  ES = 2 or 4 bytes per sample on both sides.
template <int ES>
__global__ void __launch_bounds__(256) fits_decode_kernel(const unsigned char *__restrict__ raw, unsigned char *__restrict__ dst,
		size_t bytes, int rescale) {
	const size_t head = min(bytes, (size_t) ((16 - ((uintptr_t) dst & 15)) & 15));
	const size_t nvec = (bytes - head) >> 4;
	const size_t tid = (size_t) blockIdx.x * blockDim.x + threadIdx.x, nth = (size_t) gridDim.x * blockDim.x;
	const uint4 *rv = (const uint4 *) (raw + head);
	uint4 *dv = (uint4 *) (dst + head);
	for (size_t i = tid; i < nvec; i += nth) {
		uint4 v = __ldg(rv + i);
		if (ES == 2) { v.x = dec16x2(v.x); v.y = dec16x2(v.y); v.z = dec16x2(v.z); v.w = dec16x2(v.w); }
		else { v.x = dec32(v.x, rescale); v.y = dec32(v.y, rescale); v.z = dec32(v.z, rescale); v.w = dec32(v.w, rescale); }
		dv[i] = v;
	}
	// head and tail samples (fewer than 16 bytes each)
	const size_t tail0 = head + (nvec << 4);
	const size_t nscal = (head + (bytes - tail0)) / ES;
	for (size_t i = tid; i < nscal; i += nth) {
		const size_t off = i * ES < head ? i * ES : tail0 + (i * ES - head);
		if (ES == 2) {
			const uint16_t w = *(const uint16_t *) (raw + off);
			*(uint16_t *) (dst + off) = (uint16_t) (((w >> 8) | (w << 8)) ^ 0x8000u);
		} else *(uint32_t *) (dst + off) = dec32(*(const uint32_t *) (raw + off), rescale);
	}
}

// BYTE_IMG: to[i] = (WORD) data8[i] (convert_data_ushort :396-400); rare, kept scalar
__global__ void __launch_bounds__(256) fits_decode_byte_kernel(const unsigned char *__restrict__ raw, uint16_t *__restrict__ dst, size_t n) {
	const size_t nth = (size_t) gridDim.x * blockDim.x;
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nth) dst[i] = raw[i];
}

// SER frames (io/ser.c): rows are stored top-down, samples are 8-bit or 16-bit in the byte order the header names.
// `raw` holds `rows` whole file rows; file row k of the chunk goes to block row r_first - k (memory order is bottom-up).
// mode 0: 8-bit -> WORD, unscaled (ser_manage_endianess_and_depth, ser.c:524-527); 1: 16-bit little-endian; 2: big-endian
__global__ void __launch_bounds__(256) ser_decode_kernel(const unsigned char *__restrict__ raw, uint16_t *__restrict__ dst_row0,
		int width, long rows, long r_first, int mode) {
	const size_t n = (size_t) rows * width, nth = (size_t) gridDim.x * blockDim.x;
	for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += nth) {
		const long k = (long) (i / (size_t) width); const int col = (int) (i - (size_t) k * width);
		uint16_t v;
		if (mode == 0) v = raw[i];
		else { const unsigned lo = raw[2 * i], hi = raw[2 * i + 1]; v = (uint16_t) (mode == 1 ? (lo | (hi << 8)) : (hi | (lo << 8))); }
		dst_row0[(size_t) (r_first - k) * width + col] = v;
	}
}

// ---- staging -------------------------------------------------------------------------------------
static int ensure_stage(sgpu_ctx *ctx) {
	if (ctx->h_stage[0] && ctx->h_stage[1] && ctx->d_stage[0] && ctx->d_stage[1] && ctx->stage_ev[0] && ctx->stage_ev[1]) return SGPU_OK;
	sgpu_free_stage(ctx);	// a half-made set from a failed call: start over
	for (int k = 0; k < 2; k++) {
		cudaError_t e = cudaMallocHost(&ctx->h_stage[k], SGPU_STAGE_CHUNK + 16);
		if (e == cudaSuccess) e = cudaMalloc(&ctx->d_stage[k], SGPU_STAGE_CHUNK + 16);
		if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ctx->stage_ev[k], cudaEventDisableTiming);
		if (e != cudaSuccess) { sgpu_free_stage(ctx); return sgpu_cuda_fail(ctx, e, "staging buffers of the FITS reader"); }
	}
	return SGPU_OK;
}

void sgpu_free_stage(sgpu_ctx *ctx) {
	for (int k = 0; k < 2; k++) {
		if (ctx->h_stage[k]) cudaFreeHost(ctx->h_stage[k]);
		if (ctx->d_stage[k]) cudaFree(ctx->d_stage[k]);
		if (ctx->stage_ev[k]) cudaEventDestroy(ctx->stage_ev[k]);
		ctx->h_stage[k] = ctx->d_stage[k] = nullptr; ctx->stage_ev[k] = nullptr;
	}
}

struct RawSource {	// where the raw bytes of the data unit come from
	const unsigned char *mem = nullptr;	// host memory holding the data unit, or
	int fd = -1; off_t base = 0;		// an open file and the offset of the data unit
	bool pinned = false;			// mem is page-locked: H2D straight from it
};

static bool fill_part(const RawSource &src, size_t off, char *dst, size_t bytes) {
	if (src.mem) { memcpy(dst, src.mem + off, bytes); return true; }
	size_t done = 0;
	while (done < bytes) {
		const ssize_t r = pread(src.fd, dst + done, bytes - done, src.base + (off_t) (off + done));
		if (r < 0 && errno == EINTR) continue;
		if (r <= 0) return false;
		done += (size_t) r;
	}
	return true;
}

// one staging chunk: a single thread copies out of the page cache at 4-6 GB/s, a fraction of what
// the H2D leg takes, so chunks of a few MB and more are cut into slices read side by side
static int fill_from(sgpu_ctx *ctx, const RawSource &src, size_t off, void *dst, size_t bytes) {
	static const int nthr = [] {
		const char *e = getenv("SGPU_READ_THREADS");
		int n = e ? atoi(e) : 6;
		const int hw = (int) std::thread::hardware_concurrency();
		if (hw > 0 && n > hw) n = hw;
		return n < 1 ? 1 : n;
	}();
	int parts = (int) (bytes >> 22);	// at least 4 MiB per slice
	if (parts > nthr) parts = nthr;
	bool ok = true;
	if (parts <= 1) ok = fill_part(src, off, (char *) dst, bytes);
	else {
		const size_t slice = ((bytes / parts) + 4095) & ~(size_t) 4095;
		std::vector<std::thread> th;
		std::vector<char> res(parts, 1);
		for (int t = 1; t < parts; t++) {
			const size_t o = slice * t, n = o >= bytes ? 0 : (bytes - o < slice ? bytes - o : slice);
			th.emplace_back([&, t, o, n] { if (n) res[t] = fill_part(src, off + o, (char *) dst + o, n); });
		}
		res[0] = fill_part(src, off, (char *) dst, slice < bytes ? slice : bytes);
		for (auto &x : th) x.join();
		for (char r : res) ok = ok && r;
	}
	if (!ok) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "short read of FITS data (%zu bytes at %zu)", bytes, off);
	return SGPU_OK;
}

static int decode_launch(sgpu_ctx *ctx, const unsigned char *d_raw, unsigned char *dst, size_t raw_bytes, int bitpix, int rescale) {
	const size_t work = bitpix == 8 ? raw_bytes : raw_bytes / 16 + 32;
	size_t grid = (work + 255) / 256;
	const size_t cap = (size_t) ctx->sm_count * 8;
	if (grid > cap) grid = cap;
	if (grid < 1) grid = 1;
	if (bitpix == 8) fits_decode_byte_kernel<<<(unsigned) grid, 256, 0, ctx->stream>>>(d_raw, (uint16_t *) dst, raw_bytes);
	else if (bitpix == 16) fits_decode_kernel<2><<<(unsigned) grid, 256, 0, ctx->stream>>>(d_raw, dst, raw_bytes, 0);
	else fits_decode_kernel<4><<<(unsigned) grid, 256, 0, ctx->stream>>>(d_raw, dst, raw_bytes, rescale);
	ctx->launches++;
	SGPU_CUDA(ctx, cudaGetLastError());
	return SGPU_OK;
}

// one frame: for every layer the valid row range is contiguous in the data unit and in the block
static int upload_raw(sgpu_block *b, int frame, const RawSource &src, int bitpix, int frame_height, int row0, int shifty, int rescale_arg) {
	sgpu_ctx *ctx = b->ctx;
	const sgpu_block_desc &d = b->d;
	if (frame < 0 || frame >= d.nb_frames) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d out of range", frame);
	const int want = (bitpix == -32) ? SGPU_F32 : SGPU_U16;
	if (bitpix != 8 && bitpix != 16 && bitpix != -32)
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "FITS image format %d is not supported by the device reader", bitpix);
	if (d.dtype != want) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "BITPIX %d does not match the block's sample type", bitpix);
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	int rv = ensure_stage(ctx); if (rv) return rv;
	const size_t es = b->elem(), res = (size_t) abs(bitpix) / 8;	// block / raw bytes per sample
	const size_t rowb = (size_t) d.width * es, rrow = (size_t) d.width * res;
	long lo = (long) shifty - row0; if (lo < 0) lo = 0;			// as sgpu_block_upload_frame
	long hi = (long) frame_height + shifty - row0; if (hi > d.height) hi = d.height;
	for (int l = 0; l < d.nb_layers; l++) {
		unsigned char *dst = (unsigned char *) b->dev + (((size_t) frame * d.nb_layers + l) * b->plane()) * es;
		if (hi <= lo) { SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) d.height * rowb, ctx->stream)); continue; }
		if (lo > 0) SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) lo * rowb, ctx->stream));
		if (hi < d.height) SGPU_CUDA(ctx, cudaMemsetAsync(dst + (size_t) hi * rowb, 0, (size_t) (d.height - hi) * rowb, ctx->stream));
		unsigned char *out = dst + (size_t) lo * rowb;
		size_t roff = ((size_t) l * frame_height + (size_t) (row0 + lo - shifty)) * rrow;	// into the data unit
		size_t left = (size_t) (hi - lo) * rrow;
		int rescale = rescale_arg > 0;
		if (rescale_arg < 0 && bitpix == -32) {
			// no DATAMAX key: the reference looks at three or four samples of the region it has just read
			// "and hopes for the best" (image_format_fits.c:997-1003), separately for every read
			const size_t nbdata = (size_t) (hi - lo) * d.width;
			if (nbdata > 3) {
				double data_max = 0.;
				for (size_t i = 0; i < nbdata; i += nbdata / 3) {
					unsigned char q[4];
					rv = fill_from(ctx, src, roff + i * 4, q, 4); if (rv) return rv;
					const uint32_t w = ((uint32_t) q[0] << 24) | ((uint32_t) q[1] << 16) | ((uint32_t) q[2] << 8) | q[3];
					float x; memcpy(&x, &w, 4);
					data_max = (data_max > x) ? data_max : x;	// the reference's max() macro, NaN included
				}
				rescale = data_max > 10.0;
			}
		}
		while (left) {
			// the staging copy sits at the same address modulo 16 as its destination (16-byte decode)
			const size_t pad = bitpix == 8 ? 0 : ((uintptr_t) out & 15);
			const size_t n = left < SGPU_STAGE_CHUNK ? left : SGPU_STAGE_CHUNK;	// a multiple of 16 except the last
			const int k = ctx->stage_idx; ctx->stage_idx ^= 1;
			SGPU_CUDA(ctx, cudaEventSynchronize(ctx->stage_ev[k]));	// buffer k's previous chunk has been decoded
			unsigned char *d_raw = (unsigned char *) ctx->d_stage[k] + pad;
			if (src.pinned) {
				SGPU_CUDA(ctx, cudaMemcpyAsync(d_raw, src.mem + roff, n, cudaMemcpyHostToDevice, ctx->stream));
			} else {
				unsigned char *h_raw = (unsigned char *) ctx->h_stage[k] + pad;
				rv = fill_from(ctx, src, roff, h_raw, n); if (rv) return rv;
				SGPU_CUDA(ctx, cudaMemcpyAsync(d_raw, h_raw, n, cudaMemcpyHostToDevice, ctx->stream));
			}
			rv = decode_launch(ctx, d_raw, out, n, bitpix, rescale); if (rv) return rv;
			SGPU_CUDA(ctx, cudaEventRecord(ctx->stage_ev[k], ctx->stream));
			out += n / res * es; roff += n; left -= n;
		}
	}
	return SGPU_OK;
}

int sgpu_stack_host_impl(sgpu_ctx *ctx, const sgpu_block_desc *desc, const std::function<int(sgpu_block *, int, int)> &upload_one,
		const sgpu_stack_args *a, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]);	// api.cu

// ---- primary header ------------------------------------------------------------------------------
static bool card_value(const char *card, const char *key, double *v) {
	const size_t kl = strlen(key);
	if (strncmp(card, key, kl) != 0) return false;
	for (size_t i = kl; i < 8; i++) if (card[i] != ' ') return false;
	if (card[8] != '=') return false;
	char buf[72]; memcpy(buf, card + 10, 70); buf[70] = 0;
	char *sl = strchr(buf, '/'); if (sl) *sl = 0;
	for (char *p = buf; *p; p++) if (*p == 'D' || *p == 'd') *p = 'E';	// Fortran exponents
	char *end = nullptr;
	const double x = strtod(buf, &end);
	if (end == buf) return false;
	*v = x;
	return true;
}

// a string-valued card: KEY     = 'value   '  (trailing blanks dropped)
static bool card_string(const char *card, const char *key, char out[72]) {
	const size_t kl = strlen(key);
	if (strncmp(card, key, kl) != 0) return false;
	for (size_t i = kl; i < 8; i++) if (card[i] != ' ') return false;
	if (card[8] != '=') return false;
	const char *q = (const char *) memchr(card + 10, '\'', 70);
	if (!q) return false;
	size_t n = 0;
	for (const char *p = q + 1; p < card + 80 && *p != '\'' && n < 71; p++) out[n++] = *p;
	while (n && out[n - 1] == ' ') n--;
	out[n] = 0;
	return true;
}

// One header-data unit starting at byte `off` of an open file: the image keywords into `info` (data_offset = first byte
// behind the header), whether it is an image HDU (the primary, or XTENSION = 'IMAGE'), its EXTNAME, ZIMAGE (a tile-
// compressed image stored as a binary table) and where the next HDU starts (the data unit is |BITPIX| / 8 * GCOUNT *
// (PCOUNT + NAXIS1 * ... * NAXISn) bytes, padded to 2880).  false: no HDU here (end of file, or not FITS).
struct HduScan { sgpu_fits_info info; int naxis; bool is_image, compressed; char extname[72]; long next; };
static bool scan_hdu(int fd, long off, long file_size, bool primary, HduScan *h) {
	memset(h, 0, sizeof *h);
	h->info.bscale = 1.0; h->naxis = -1;
	h->info.naxes[0] = h->info.naxes[1] = h->info.naxes[2] = 1;
	long axes[10]; for (long &a : axes) a = 1;
	long pcount = 0, gcount = 1;
	char blk[2880];
	bool first = true, done = false;
	while (!done) {
		size_t got = 0;
		while (got < sizeof blk) {
			const ssize_t r = pread(fd, blk + got, sizeof blk - got, off + (long) got);
			if (r < 0 && errno == EINTR) continue;
			if (r <= 0) break;
			got += (size_t) r;
		}
		if (got < sizeof blk) return false;
		if (first) {
			char xt[72];
			if (primary) { if (strncmp(blk, "SIMPLE  =", 9) != 0) return false; h->is_image = true; }
			else { if (!card_string(blk, "XTENSION", xt)) return false; h->is_image = strcmp(xt, "IMAGE") == 0; }
		}
		first = false;
		off += (long) sizeof blk;
		for (int c = 0; c < 36; c++) {
			const char *card = blk + 80 * c;
			double v; char sv[72];
			if (strncmp(card, "END     ", 8) == 0) { done = true; break; }
			if (card_value(card, "BITPIX", &v)) h->info.bitpix = (int) v;
			else if (card_value(card, "NAXIS", &v)) h->naxis = (int) v;
			else if (strncmp(card, "NAXIS", 5) == 0 && card[5] >= '1' && card[5] <= '9' && card[6] == ' ') {
				char key[8] = "NAXIS0"; key[5] = card[5];
				if (card_value(card, key, &v)) { axes[card[5] - '0'] = (long) v; if (card[5] <= '3') h->info.naxes[card[5] - '1'] = (long) v; }
			}
			else if (card_value(card, "PCOUNT", &v)) pcount = (long) v;
			else if (card_value(card, "GCOUNT", &v)) gcount = (long) v;
			else if (card_value(card, "BZERO", &v)) h->info.bzero = v;
			else if (card_value(card, "BSCALE", &v)) h->info.bscale = v;
			else if (card_value(card, "DATAMAX", &v)) { h->info.datamax = v; h->info.has_datamax = 1; }
			else if (card_string(card, "EXTNAME", sv)) memcpy(h->extname, sv, sizeof sv);
			else if (strncmp(card, "ZIMAGE  =", 9) == 0 && memchr(card + 10, 'T', 21)) h->compressed = true;
		}
	}
	if (h->naxis < 0 || h->naxis > 9 || h->info.bitpix == 0) return false;
	long n = h->naxis ? 1 : 0;
	for (int k = 1; k <= h->naxis; k++) n *= axes[k];
	const long bytes = (labs((long) h->info.bitpix) / 8) * gcount * (pcount + n);
	h->info.data_offset = off;
	h->next = off + (bytes + 2879) / 2880 * 2880;
	if (off + bytes > file_size) return false;
	const bool plain = h->info.bscale == 1.0;
	h->info.device_readable = h->is_image && !h->compressed && plain && ((h->info.bitpix == 8 && h->info.bzero == 0.0) ||
			(h->info.bitpix == 16 && (h->info.bzero == 0.0 || h->info.bzero == 32768.0)) ||
			(h->info.bitpix == -32 && h->info.bzero == 0.0));
	return true;
}

extern "C" {

/* The frames of a FITS sequence file (SEQ_FITSEQ: one image per HDU), chosen as _find_hdus does (io/fits_sequence.c:39-150):
 * image HDUs with NAXIS > 1, not named ICCProfile / Thumbnail / HDRLET / WCS / D2IM; an HDU whose layer count or sample
 * type differs from the first image's is skipped, one whose width or height differs fails the file. */
int sgpu_fitseq_probe(const char *path, sgpu_fits_info *infos, int max_frames, int *nb_frames) {
	if (!path || !nb_frames) return SGPU_GENERIC_ERROR;
	*nb_frames = 0;
	const int fd = open(path, O_RDONLY);
	if (fd < 0) return SGPU_SEQUENCE_ERROR;
	struct stat st;
	if (fstat(fd, &st) != 0) { close(fd); return SGPU_SEQUENCE_ERROR; }
	long off = 0; bool primary = true, have_ref = false; int rv = SGPU_OK, count = 0;
	sgpu_fits_info ref; memset(&ref, 0, sizeof ref);
	auto kind = [](const sgpu_fits_info &i) { return i.bitpix == 16 && i.bzero == 32768.0 ? 20 : i.bitpix; };	// manage_bitpix: USHORT_IMG
	while (off < (long) st.st_size) {
		HduScan h;
		if (!scan_hdu(fd, off, (long) st.st_size, primary, &h)) { if (primary) rv = SGPU_SEQUENCE_ERROR; break; }
		primary = false; off = h.next;
		if (!h.is_image) continue;
		static const char *skip[] = { "ICCProfile", "Thumbnail", "HDRLET", "WCS", "D2IM" };
		bool skipped = false;
		for (const char *sname : skip) if (strncmp(h.extname, sname, strlen(sname)) == 0) skipped = true;
		if (skipped || h.naxis <= 1) continue;
		if (!have_ref) { ref = h.info; have_ref = true; }
		else {
			if (h.info.naxes[2] != ref.naxes[2] || kind(h.info) != kind(ref)) continue;
			if (h.info.naxes[0] != ref.naxes[0] || h.info.naxes[1] != ref.naxes[1]) { rv = SGPU_SEQUENCE_ERROR; break; }	// "different parameters, which is not allowed"
		}
		if (infos && count < max_frames) infos[count] = h.info;
		count++;
	}
	close(fd);
	if (rv) return rv;
	*nb_frames = count;
	return SGPU_OK;
}

int sgpu_fits_probe(const char *path, sgpu_fits_info *info) {
	if (!path || !info) return SGPU_GENERIC_ERROR;
	memset(info, 0, sizeof *info);
	info->bscale = 1.0;
	const int fd = open(path, O_RDONLY);
	if (fd < 0) return SGPU_SEQUENCE_ERROR;
	char blk[2880];
	long off = 0;
	int naxis = -1, rv = SGPU_SEQUENCE_ERROR;
	bool first = true, done = false;
	info->naxes[0] = info->naxes[1] = info->naxes[2] = 1;
	while (!done) {
		size_t got = 0;
		while (got < sizeof blk) {
			const ssize_t r = pread(fd, blk + got, sizeof blk - got, off + (long) got);
			if (r < 0 && errno == EINTR) continue;
			if (r <= 0) break;
			got += (size_t) r;
		}
		if (got < sizeof blk) break;
		if (first && strncmp(blk, "SIMPLE  =", 9) != 0) break;
		first = false;
		off += (long) sizeof blk;
		for (int c = 0; c < 36; c++) {
			const char *card = blk + 80 * c;
			double v;
			if (strncmp(card, "END     ", 8) == 0) { done = true; break; }
			if (card_value(card, "BITPIX", &v)) info->bitpix = (int) v;
			else if (card_value(card, "NAXIS", &v)) naxis = (int) v;
			else if (card_value(card, "NAXIS1", &v)) info->naxes[0] = (long) v;
			else if (card_value(card, "NAXIS2", &v)) info->naxes[1] = (long) v;
			else if (card_value(card, "NAXIS3", &v)) info->naxes[2] = (long) v;
			else if (card_value(card, "BZERO", &v)) info->bzero = v;
			else if (card_value(card, "BSCALE", &v)) info->bscale = v;
			else if (card_value(card, "DATAMAX", &v)) { info->datamax = v; info->has_datamax = 1; }
		}
	}
	struct stat st;
	if (done && naxis >= 2 && naxis <= 3 && info->bitpix != 0 && fstat(fd, &st) == 0) {
		info->data_offset = off;
		const long need = info->naxes[0] * info->naxes[1] * info->naxes[2] * (labs((long) info->bitpix) / 8);
		if (info->naxes[0] > 0 && info->naxes[1] > 0 && info->naxes[2] > 0 && st.st_size >= off + need) rv = SGPU_OK;
	}
	close(fd);
	if (rv) return rv;
	// what the device reader decodes without cfitsio: BYTE, SHORT (BZERO 0), USHORT (BZERO 32768) and
	// FLOAT without physical scaling (fits_uses_physical_scaling, image_format_fits.c:521-553)
	const bool plain = info->bscale == 1.0;
	info->device_readable = plain && ((info->bitpix == 8 && info->bzero == 0.0) ||
			(info->bitpix == 16 && (info->bzero == 0.0 || info->bzero == 32768.0)) ||
			(info->bitpix == -32 && info->bzero == 0.0));
	return SGPU_OK;
}

int sgpu_block_upload_fits(sgpu_block *b, int frame, const void *data_unit, int bitpix,
		int frame_height, int row0, int shifty, int rescale_ushrt) {
	if (!b || !data_unit) return SGPU_GENERIC_ERROR;
	RawSource src; src.mem = (const unsigned char *) data_unit;
	cudaPointerAttributes at;
	if (cudaPointerGetAttributes(&at, data_unit) == cudaSuccess) src.pinned = at.type == cudaMemoryTypeHost;
	else cudaGetLastError();
	return upload_raw(b, frame, src, bitpix, frame_height, row0, shifty, rescale_ushrt);
}

int sgpu_block_upload_fits_fd(sgpu_block *b, int frame, int fd, long data_offset, int bitpix,
		int frame_height, int row0, int shifty, int rescale_ushrt) {
	if (!b || fd < 0 || data_offset < 0) return SGPU_GENERIC_ERROR;
	RawSource src; src.fd = fd; src.base = (off_t) data_offset;
	return upload_raw(b, frame, src, bitpix, frame_height, row0, shifty, rescale_ushrt);
}

// the block loop over frames that are data units of open files: fds[f] / fi[f] (several frames may share a descriptor)
static int stack_fits_units(sgpu_ctx *ctx, int nb_frames, const std::vector<int> &fds, const std::vector<sgpu_fits_info> &fi, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]) {
	for (int f = 0; f < nb_frames; f++) {
		if (!fi[f].device_readable) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "image %d (BITPIX %d, BZERO %g, BSCALE %g) needs cfitsio", f, fi[f].bitpix, fi[f].bzero, fi[f].bscale);
		// "stacking an image sequence with different image sizes is not possible" (median_and_mean.c:1300-1310)
		if (fi[f].bitpix != fi[0].bitpix || fi[f].naxes[0] != fi[0].naxes[0] || fi[f].naxes[1] != fi[0].naxes[1] || fi[f].naxes[2] != fi[0].naxes[2])
			return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "image %d differs in size or sample type from the first image", f);
	}
	if (fi[0].naxes[2] > 3) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "more than three layers");
	const sgpu_block_desc desc = { nb_frames, (int) fi[0].naxes[2], (int) fi[0].naxes[0], (int) fi[0].naxes[1],
			fi[0].bitpix == -32 ? SGPU_F32 : SGPU_U16 };
	const int H = desc.height;
	// Float data without a DATAMAX key: the reference guesses the range from three or four samples of
	// whatever region it has just read (image_format_fits.c:997-1003), so its answer depends on ITS
	// block plan.  The row blocks here are not the reference's, and a frame must not be scaled in one
	// block and left alone in another: the guess is made ONCE per frame, on the samples the reference
	// would look at when it reads the first layer whole.
	std::vector<int> rescale(nb_frames, 0);
	for (int f = 0; f < nb_frames; f++) {
		if (fi[f].bitpix != -32) continue;
		if (fi[f].has_datamax) { rescale[f] = fi[f].datamax > 10.0; continue; }
		const size_t nbdata = (size_t) desc.width * (size_t) H;
		if (nbdata <= 3) continue;
		double data_max = 0.;
		for (size_t i = 0; i < nbdata; i += nbdata / 3) {
			unsigned char q[4];
			if (pread(fds[f], q, 4, (off_t) (fi[f].data_offset + i * 4)) != 4) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot read image %d", f);
			const uint32_t w = ((uint32_t) q[0] << 24) | ((uint32_t) q[1] << 16) | ((uint32_t) q[2] << 8) | q[3];
			float x; memcpy(&x, &w, 4);
			data_max = (data_max > x) ? data_max : x;	// the reference's max() macro, NaN included
		}
		rescale[f] = data_max > 10.0;
	}
	sgpu_stack_args with_bitpix = *args;
	with_bitpix.source_bitpix = fi[0].bitpix;	// BYTE_IMG sequences: normalize_to16bit
	return sgpu_stack_host_impl(ctx, &desc, [&](sgpu_block *b, int f, int row0) {
		RawSource src; src.fd = fds[f]; src.base = (off_t) fi[f].data_offset;
		return upload_raw(b, f, src, fi[f].bitpix, H, row0, shifty ? shifty[f] : 0, rescale[f]);
	}, &with_bitpix, out, rej_low, rej_high, irej);
}

int sgpu_stack_fits(sgpu_ctx *ctx, int nb_frames, const char *const *paths, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]) {
	if (!ctx || !paths || nb_frames < 1 || !out) return SGPU_GENERIC_ERROR;
	std::vector<sgpu_fits_info> fi(nb_frames);
	std::vector<int> fds(nb_frames, -1);
	auto close_all = [&](int code) { for (int fd : fds) if (fd >= 0) close(fd); return code; };
	for (int f = 0; f < nb_frames; f++) {
		if (!paths[f] || sgpu_fits_probe(paths[f], &fi[f])) return close_all(sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot read image %d of the sequence as FITS", f));
		fds[f] = open(paths[f], O_RDONLY);
		if (fds[f] < 0) return close_all(sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot open image %d", f));
	}
	return close_all(stack_fits_units(ctx, nb_frames, fds, fi, shifty, args, out, rej_low, rej_high, irej));
}

/* sgpu_stack_fits for a FITS sequence file: frame_indices (or NULL: all, in file order) select among the frames
 * sgpu_fitseq_probe finds (args->image_indices over fitseq->hdu_index, io/fits_sequence.c:285-300) */
int sgpu_stack_fitseq(sgpu_ctx *ctx, const char *path, int nb_frames, const int *frame_indices, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]) {
	if (!ctx || !path || nb_frames < 1 || !out) return SGPU_GENERIC_ERROR;
	int total = 0;
	if (sgpu_fitseq_probe(path, nullptr, 0, &total) || total < 1) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot read %s as a FITS sequence", path);
	std::vector<sgpu_fits_info> all(total), fi(nb_frames);
	if (sgpu_fitseq_probe(path, all.data(), total, &total)) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot read %s as a FITS sequence", path);
	for (int f = 0; f < nb_frames; f++) {
		const int k = frame_indices ? frame_indices[f] : f;
		if (k < 0 || k >= total) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d is not in the sequence file (%d frames)", k, total);
		fi[f] = all[k];
	}
	const int fd = open(path, O_RDONLY);
	if (fd < 0) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot open %s", path);
	std::vector<int> fds(nb_frames, fd);
	const int rv = stack_fits_units(ctx, nb_frames, fds, fi, shifty, args, out, rej_low, rej_high, irej);
	close(fd);
	return rv;
}

/* ---- SER files (io/ser.c) ---- */
int sgpu_ser_probe(const char *path, sgpu_ser_info *info) {
	if (!path || !info) return SGPU_GENERIC_ERROR;
	memset(info, 0, sizeof *info);
	const int fd = open(path, O_RDONLY);
	if (fd < 0) return SGPU_SEQUENCE_ERROR;
	unsigned char h[178];
	struct stat st;
	const bool ok = pread(fd, h, sizeof h, 0) == (ssize_t) sizeof h && fstat(fd, &st) == 0;
	close(fd);
	if (!ok) return SGPU_SEQUENCE_ERROR;
	auto le32 = [&](int off) { return (int) ((uint32_t) h[off] | ((uint32_t) h[off + 1] << 8) | ((uint32_t) h[off + 2] << 16) | ((uint32_t) h[off + 3] << 24)); };
	info->color_id = le32(18); info->little_endian = le32(22); info->width = le32(26); info->height = le32(30);
	info->bit_depth = le32(34); info->frame_count = le32(38);
	info->byte_depth = info->bit_depth <= 8 ? 1 : 2;				// ser.c:345-347
	info->planes = (info->color_id == 100 || info->color_id == 101) ? 3 : 1;	// SER_RGB / SER_BGR (:349-352)
	info->data_offset = 178;
	if (info->width < 1 || info->height < 1 || info->frame_count < 1 || info->bit_depth < 1 || info->bit_depth > 16) return SGPU_SEQUENCE_ERROR;
	const long long frame = (long long) info->width * info->height * info->planes * info->byte_depth;
	if ((long long) st.st_size < 178 + frame * info->frame_count) return SGPU_SEQUENCE_ERROR;
	// one plane per frame read as it is (mono, or a colour filter array taken as mono: no demosaicing here)
	info->device_readable = info->planes == 1;
	return SGPU_OK;
}

int sgpu_block_upload_ser_fd(sgpu_block *b, int frame, int fd, const sgpu_ser_info *info, int ser_frame, int row0, int shifty) {
	if (!b || fd < 0 || !info) return SGPU_GENERIC_ERROR;
	sgpu_ctx *ctx = b->ctx;
	const sgpu_block_desc &d = b->d;
	if (frame < 0 || frame >= d.nb_frames) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d out of range", frame);
	if (ser_frame < 0 || ser_frame >= info->frame_count) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "SER frame %d out of range", ser_frame);
	if (!info->device_readable || d.nb_layers != 1 || d.dtype != SGPU_U16 || d.width != info->width)
		return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "this SER file does not fit the block (one 16-bit layer, %d wide)", d.width);
	SGPU_CUDA(ctx, cudaSetDevice(ctx->device));
	int rv = ensure_stage(ctx); if (rv) return rv;
	const int H = info->height, W = info->width;
	const size_t bpp = (size_t) info->byte_depth, rrow = (size_t) W * bpp, rowb = (size_t) W * 2;
	long lo = (long) shifty - row0; if (lo < 0) lo = 0;		// as sgpu_block_upload_frame, in memory (bottom-up) rows
	long hi = (long) H + shifty - row0; if (hi > d.height) hi = d.height;
	unsigned char *dst = (unsigned char *) b->dev + ((size_t) frame * b->plane()) * 2;
	if (hi <= lo) { SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) d.height * rowb, ctx->stream)); return SGPU_OK; }
	if (lo > 0) SGPU_CUDA(ctx, cudaMemsetAsync(dst, 0, (size_t) lo * rowb, ctx->stream));
	if (hi < d.height) SGPU_CUDA(ctx, cudaMemsetAsync(dst + (size_t) hi * rowb, 0, (size_t) (d.height - hi) * rowb, ctx->stream));
	// block row r holds memory row m = row0 + r - shifty of the frame = FILE row H - 1 - m (SER rows are top-down): block rows
	// [lo, hi) are the file rows [f0, f0 + hi - lo) read forwards and written backwards from block row hi - 1
	const long f0 = (long) H - 1 - ((long) row0 + hi - 1 - shifty);
	const int mode = info->byte_depth == 1 ? 0 : (info->little_endian == 1 ? 2 : 1);	// header value 1 = big-endian DATA (ser.h:71-80)
	const off_t base = (off_t) info->data_offset + (off_t) ((size_t) H * rrow) * ser_frame + (off_t) f0 * (off_t) rrow;
	RawSource src; src.fd = fd; src.base = base;
	const long rows_per_chunk = (long) (SGPU_STAGE_CHUNK / rrow) > 0 ? (long) (SGPU_STAGE_CHUNK / rrow) : 1;
	if (rrow > SGPU_STAGE_CHUNK) return sgpu_fail(ctx, SGPU_GENERIC_ERROR, "SER rows of %zu bytes exceed the staging chunk", rrow);
	for (long done = 0; done < hi - lo; done += rows_per_chunk) {
		const long nr = (hi - lo - done) < rows_per_chunk ? (hi - lo - done) : rows_per_chunk;
		const size_t n = (size_t) nr * rrow;
		const int k = ctx->stage_idx; ctx->stage_idx ^= 1;
		SGPU_CUDA(ctx, cudaEventSynchronize(ctx->stage_ev[k]));
		rv = fill_from(ctx, src, (size_t) done * rrow, ctx->h_stage[k], n); if (rv) return rv;
		SGPU_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage[k], ctx->h_stage[k], n, cudaMemcpyHostToDevice, ctx->stream));
		size_t grid = ((size_t) nr * W + 255) / 256; const size_t cap = (size_t) ctx->sm_count * 8;
		if (grid > cap) grid = cap;
		ser_decode_kernel<<<(unsigned) grid, 256, 0, ctx->stream>>>((const unsigned char *) ctx->d_stage[k], (uint16_t *) dst, W, nr, hi - 1 - done, mode);
		ctx->launches++;
		SGPU_CUDA(ctx, cudaGetLastError());
		SGPU_CUDA(ctx, cudaEventRecord(ctx->stage_ev[k], ctx->stream));
	}
	return SGPU_OK;
}

int sgpu_stack_ser(sgpu_ctx *ctx, const char *path, int nb_frames, const int *frame_indices, const int *shifty,
		const sgpu_stack_args *args, void *out, uint16_t *rej_low, uint16_t *rej_high, uint64_t irej[6]) {
	if (!ctx || !path || nb_frames < 1 || !out) return SGPU_GENERIC_ERROR;
	sgpu_ser_info si;
	if (sgpu_ser_probe(path, &si)) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot read %s as a SER file", path);
	if (!si.device_readable) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "%s holds colour frames (ColorID %d): Siril's reader", path, si.color_id);
	for (int f = 0; f < nb_frames; f++) {
		const int k = frame_indices ? frame_indices[f] : f;
		if (k < 0 || k >= si.frame_count) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "frame %d is not in the SER file (%d frames)", k, si.frame_count);
	}
	const int fd = open(path, O_RDONLY);
	if (fd < 0) return sgpu_fail(ctx, SGPU_SEQUENCE_ERROR, "cannot open %s", path);
	const sgpu_block_desc desc = { nb_frames, 1, si.width, si.height, SGPU_U16 };
	sgpu_stack_args with_bitpix = *args;
	with_bitpix.source_bitpix = si.byte_depth == 1 ? 8 : 16;	// stack_open_all_files: BYTE_IMG or USHORT_IMG (median_and_mean.c:220)
	const int rv = sgpu_stack_host_impl(ctx, &desc, [&](sgpu_block *b, int f, int row0) {
		return sgpu_block_upload_ser_fd(b, f, fd, &si, frame_indices ? frame_indices[f] : f, row0, shifty ? shifty[f] : 0);
	}, &with_bitpix, out, rej_low, rej_high, irej);
	close(fd);
	return rv;
}

}	// extern "C"
