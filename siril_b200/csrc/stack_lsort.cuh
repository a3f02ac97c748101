// stack_lsort.cuh — kernel of the LANE-SORT engine: same coverage as the sorted engine
// (stack_sorted.cuh) with a sort that needs an order of magnitude fewer shuffles and a rejection
// phase that keeps every thread busy.
//
// One CTA of SUB warps owns SUB sub-tiles of 32 consecutive pixels of one layer (tile load: TMA, one
// 3-D box per sub-tile and <= 256 frames, 128B swizzle; or coalesced row loads with x shifts /
// 16-bit samples — stack_sorted.cuh's loaders).
//  sort    a group of T = 2^LT lanes owns one pixel; lane t takes the samples of rows j*T + t
//          (bank-conflict free) into RE registers (+inf beyond the stack and for nulls in mean
//          mode).  Each lane first sorts its own run with a generated odd-even merge network
//          (gen_sortnet.py; only the RS rows that can hold samples take part), then LT merge
//          levels combine the runs LANE-MAJOR (lane t ends up with ranks t*RE .. t*RE+RE-1): per
//          level one flip exchange (run j against the partner's run RE-1-j), L-1 lane half-cleaners
//          and a register-only bitonic merge of the lane's run.  Shuffles per sample: LT(LT+1)/2
//          instead of ~log2(N)^2/2.  The group leaves n and sum(x-K), sum((x-K)^2) beside the stack.
//  reject  after all sort rounds, thread i of the CTA runs the per-pixel rejection / combine
//          (sorted.cuh) on pixel i of the tile: SUB*32 threads, SUB*32 pixels.
#pragma once
#include "stack_sorted.cuh"

namespace sgpu {

template <int RS> struct LocalNet { template <int RE> static __device__ __forceinline__ void run(float (&v)[RE]); };
#define LCE(a, b) { const float lo__ = fminf(v[a], v[b]); v[b] = fmaxf(v[a], v[b]); v[a] = lo__; }
#define SGPU_LOCALNET(RS_) template <> template <int RE> __device__ __forceinline__ void LocalNet<RS_>::run(float (&v)[RE])

// register-only merge of a bitonic run of RE values, ascending.  A half-cleaner (i against
// i + LEN/2) splits a bitonic run of any EVEN length into two bitonic halves with every value of
// the first <= every value of the second; an odd tail of 3 or 5 values is sorted outright.  So RE
// may be 2^k, 3 * 2^k or 5 * 2^k (20 and 24 slots serve stacks that would waste a third of 32).
#define LCE2(a, b) { const float lo__ = fminf(v[a], v[b]); v[b] = fmaxf(v[a], v[b]); v[a] = lo__; }
template <int LEN, int OFF, int RE>
__device__ __forceinline__ void merge_bitonic_run(float (&v)[RE]) {
	if constexpr (LEN <= 1) return;
	else if constexpr (LEN % 2 == 0) {
#pragma unroll
		for (int i = 0; i < LEN / 2; i++) LCE2(OFF + i, OFF + i + LEN / 2)
		merge_bitonic_run<LEN / 2, OFF, RE>(v);
		merge_bitonic_run<LEN / 2, OFF + LEN / 2, RE>(v);
	} else if constexpr (LEN == 3) {
		LCE2(OFF, OFF + 1) LCE2(OFF + 1, OFF + 2) LCE2(OFF, OFF + 1)
	} else {
		static_assert(LEN == 5, "run length must be 2^k, 3 * 2^k or 5 * 2^k");
		LCE2(OFF, OFF + 1) LCE2(OFF + 3, OFF + 4) LCE2(OFF + 2, OFF + 4) LCE2(OFF + 2, OFF + 3) LCE2(OFF + 1, OFF + 4)
		LCE2(OFF, OFF + 3) LCE2(OFF, OFF + 2) LCE2(OFF + 1, OFF + 3) LCE2(OFF + 1, OFF + 2)
	}
}
template <int RE>
__device__ __forceinline__ void local_bitonic_merge(float (&v)[RE]) { merge_bitonic_run<RE, 0, RE>(v); }

// merge the T sorted runs of a group lane-major: afterwards lane t holds ranks [t*RE, (t+1)*RE)
template <int LT, int RE>
__device__ __forceinline__ void group_merge(float (&v)[RE], int t, unsigned mask) {
#pragma unroll
	for (int L = 1; L <= LT; L++) {
		{	// flip: element j of this run against element RE-1-j of the partner's
			const int pm = (1 << L) - 1;
			const bool keep_min = ((t >> (L - 1)) & 1) == 0;
#pragma unroll
			for (int j = 0; j < RE / 2; j++) {
				const float a = v[j], b = v[RE - 1 - j];
				const float xa = __shfl_xor_sync(mask, b, pm), xb = __shfl_xor_sync(mask, a, pm);
				v[j] = keep_min ? fminf(a, xa) : fmaxf(a, xa);
				v[RE - 1 - j] = keep_min ? fminf(b, xb) : fmaxf(b, xb);
			}
		}
#pragma unroll
		for (int s = 1 << (L - 2 >= 0 ? L - 2 : 0); L >= 2 && s > 0; s >>= 1) {	// half-cleaners between lanes
			const bool keep_min = (t & s) == 0;
#pragma unroll
			for (int j = 0; j < RE; j++) {
				const float x = __shfl_xor_sync(mask, v[j], s);
				v[j] = keep_min ? fminf(v[j], x) : fmaxf(v[j], x);
			}
		}
		local_bitonic_merge<RE>(v);
	}
}

// sorted stack of one pixel in the tile: rank i lives in row i
struct SubStack {
	char *base; int c;	// base: sub-tile + (px & 3) * 4 bytes; c = px >> 2
	__device__ __forceinline__ SubStack(float *rows, int px) : base((char *) rows + ((px & 3) << 2)), c(px >> 2) { }
	__device__ __forceinline__ float *at(int i) const { return (float *) (base + ((((i << 3) | ((i ^ c) & 7))) << 4)); }
	__device__ __forceinline__ float operator()(int i) const { return *at(i); }
	__device__ __forceinline__ void set(int i, float v) const { *at(i) = v; }
};
// ranks [i0, i0 + 8), i0 a multiple of 8: one base address, eight fixed offsets (the swizzle term of
// row i0 + k is (k ^ c) & 7)
__device__ __forceinline__ void acc_load8(const SubStack &s, int i0, int n, float (&x)[8]) {
	const char *q = s.base + (size_t) i0 * 128;
#pragma unroll
	for (int k = 0; k < 8; k++) x[k] = (i0 + k < n) ? *(const float *) (q + k * 128 + (((k ^ s.c) & 7) << 4)) : 0.f;
}

// pixel (within a sub-tile) of group g of a warp in sort round r (T rounds cover the 32 pixels):
// chosen so that the lanes of a warp hit different banks when they read rows j*T + t of the
// 128B-swizzled tile (bank = (((px >> 2) ^ (row & 7)) << 2) | (px & 3))
template <int LT>
__device__ __forceinline__ int lsort_pixel(int r, int g) {
	if (LT == 1) return ((r ^ (2 * (g >> 2))) << 2) | (g & 3);	// 16 px / round: 16-byte chunks r, r^2, r^4, r^6
	if (LT == 2) return ((r + 4 * (g >> 2)) << 2) | (g & 3);	// 8 px / round: chunks r and r + 4
	if (LT == 3) return (r << 2) | g;				// 4 px / round: chunk r
	return r * (32 >> LT) + g;					// 2 or 1 px / round
}


// ---- GESDT with the lanes of a small group (sorted.cuh: fast_gesdt is the one-thread form and says what is computed) ----
// Only the DECISION of an iteration — which end of the window goes — is serial, and it needs no more than the window's sum:
// one thread per pixel walks it (gesdt_decide; every lane of the warp busy) and leaves one bit per iteration in shared
// memory.  The Grubbs tests of all iterations are then evaluated side by side by GL lanes per pixel (coop_gesdt): every
// lane takes a contiguous chunk of iterations, sums what its chunk removes, an exclusive scan over the lanes gives each the
// sums its chunk starts from, and it replays the tests of its chunk from there.  The last iteration that certainly beat its
// critical value (and the last too close to call) is a maximum over the lanes; lane 0 then collects what the reference
// keeps up to that iteration.  The one-thread form is a chain of ~100 dependent instructions per iteration walked by one
// warp per CTA; here the serial part is a chain of a dozen.
#define SGPU_GESDT_COOP_MAX 128
#define SGPU_GESDT_WORDS (SGPU_GESDT_COOP_MAX / 32 + 1)	// per pixel in shared memory: the decision bits, then a flag (the float mean sat on a rounding boundary)
template <bool U16>
__device__ __forceinline__ bool gesdt_exact_sums(const SubStack &s, int n) { return U16 || (s(0) > 0.f && s(n - 1) <= s(0) * 65536.f && n <= 2048); }

// word w of the pixel goes to out[w * stride]
template <bool U16>
__device__ __forceinline__ void gesdt_decide(const SubStack &s, int n, int max_out, const RejCfg &c, const Sums &S, unsigned *out, int stride) {
	const bool exact_sums = gesdt_exact_sums<U16>(s, n);
	int fail = 0;
	double a = S.S1 + (double) n * S.K;	// the window's sum; its roundings are far inside the room `tiny` leaves
	int lo = 0, size = n;
#pragma unroll
	for (int w = 0; w < SGPU_GESDT_COOP_MAX / 32; w++) {
		unsigned word = 0;
		const int lim = min(32, max_out - 32 * w);
		for (int j = 0; j < lim; j++, size--) {
			const float xl = s(lo), xh = s(lo + size - 1);
			const double r0 = rcp_t(c.rcp, size), b = (double) size, al = a - (double) xl, ah = a - (double) xh;
			double avg_d = a * r0;
			avg_d = fma(fma(-avg_d, b, a), r0, avg_d);
			const float avg = (float) avg_d;
			if (!exact_sums) {
				const double tiny = fabs(avg_d) * 1e-12;
				if ((float) (avg_d + tiny) != avg || (float) (avg_d - tiny) != avg) fail = 1;
			}
			const bool low_end = !(xh - avg > avg - xl);
			word |= (unsigned) low_end << j;
			lo += low_end;
			a = low_end ? al : ah;
		}
		out[w * stride] = word;
	}
	out[(SGPU_GESDT_COOP_MAX / 32) * stride] = (unsigned) fail;
}

// `active` false: the lane only takes part in the shuffles
template <bool U16, int GL>
__device__ __forceinline__ void coop_gesdt(const SubStack &s, int n, int removed, int max_out, const RejCfg &c, const Sums &S, int q, bool active,
		const unsigned *dec, int stride, int rej[2], FastRes &o) {
	o.status = FAST_FALLBACK;
	double median = 0.0;
	unsigned bits[SGPU_GESDT_COOP_MAX / 32] = { 0, 0, 0, 0 };	// (every index below is a compile-time one: the words stay in registers)
	int fail = 0;
	if (active) {
		if (n & 1) median = (double) s(n / 2);
		else if (U16) median = ((double) s((n - 1) / 2) + (double) s(n / 2)) / 2.0;
		else median = (s((n - 1) / 2) + s(n / 2)) / 2.0;
		if (U16) median = (double) (float) median;
#pragma unroll
		for (int w = 0; w < SGPU_GESDT_COOP_MAX / 32; w++) bits[w] = dec[w * stride];
		fail = (int) dec[(SGPU_GESDT_COOP_MAX / 32) * stride];
	}
	const double K = S.K;
	auto bit = [&](int it) {
		const unsigned w01 = it < 32 ? bits[0] : bits[1], w23 = it < 96 ? bits[2] : bits[3];
		return ((it < 64 ? w01 : w23) >> (it & 31)) & 1u;
	};
	auto lows_before = [&](int it) {	// low-end removals among iterations [0, it)
		int k = 0;
#pragma unroll
		for (int w = 0; w < SGPU_GESDT_COOP_MAX / 32; w++) {
			const int left = it - 32 * w;
			if (left > 0) k += __popc(left >= 32 ? bits[w] : (bits[w] & ((1u << left) - 1u)));
		}
		return k;
	};
	// ---- the tests, a chunk of iterations per lane ----
	const int ch = (max_out + GL - 1) / GL;
	const int it0 = active ? min(q * ch, max_out) : 0, it1 = active ? min(it0 + ch, max_out) : 0;
	double A1 = 0.0, A2 = 0.0;
	{
		int lo = lows_before(it0);
		for (int it = it0; it < it1; it++) {
			const bool low = bit(it);
			const float xrem = low ? s(lo) : s(n - 1 - (it - lo));
			lo += low;
			const double y = (double) xrem - K;
			A1 += y; A2 = fma(y, y, A2);
		}
	}
	double P1 = A1, P2 = A2;	// inclusive scan over the group's lanes, then made exclusive
#pragma unroll
	for (int d = 1; d < GL; d <<= 1) {
		const double u1 = __shfl_up_sync(0xffffffffu, P1, d, GL), u2 = __shfl_up_sync(0xffffffffu, P2, d, GL);
		if (q >= d) { P1 += u1; P2 += u2; }
	}
	double S1 = S.S1 - (P1 - A1), S2 = S.S2 - (P2 - A2);
	int last = -1, amb = -1;
	bool bad = false;
	{
		int lo = lows_before(it0);
		for (int it = it0; it < it1; it++) {
			const int size = n - it;
			const bool low = bit(it);
			const float xrem = low ? s(lo) : s(n - 1 - (it - lo));
			if (it > 1) {
				const double r0 = rcp_t(c.rcp, size), r1 = rcp_t(c.rcp, size - 1), b = (double) size, a = S1 + b * K;
				double avg_d = a * r0;
				avg_d = fma(fma(-avg_d, b, a), r0, avg_d);
				const float avg = (float) avg_d;
				const float dev = low ? avg - xrem : xrem - avg;
				const double var = (S2 - S1 * S1 * r0) * r1;
				if (!(var >= 0.0)) { bad = true; break; }
				const double e = fabs(avg_d) * 5.9604645e-08;
				const double var_hi = var + e * e * 2.0;
				const double cr = (double) c.crit[it + removed], d2 = (double) dev * (double) dev, c2 = cr * cr;
				const bool normal = var > 1e-30 && var_hi < 1e30;
				// (the sums of a chunk come from a scan, not from the serial down-dating: their last bits differ, which the
				// room left around the float path's roundings covers many times over)
				if (normal && d2 > c2 * var_hi * (1.0 + 4.0e-6)) last = it;
				else if (!normal || !(d2 < c2 * var * (1.0 - 4.0e-6))) {
					float slo, shi;
					if (!sd_interval(var, avg_d, size, SGPU_DELTA_SD, 1.f, &slo, &shi)) { bad = true; break; }
					const float crf = c.crit[it + removed];
					if (dev / shi > crf) last = it;
					else if (dev / slo > crf) amb = it;
				}
			}
			lo += low;
			const double y = (double) xrem - K;
			S1 -= y; S2 = fma(-y, y, S2);
		}
	}
#pragma unroll
	for (int d = 1; d < GL; d <<= 1) {
		last = max(last, __shfl_xor_sync(0xffffffffu, last, d, GL));
		amb = max(amb, __shfl_xor_sync(0xffffffffu, amb, d, GL));
		bad = __shfl_xor_sync(0xffffffffu, (int) bad, d, GL) || bad;
	}
	if (!active || q != 0 || fail || bad) return;
	last = max(last, min(1, max_out - 1));	// iterations 0 and 1 are always taken (confirm_outliers never keeps fewer than two)
	if (amb > last) return;
	// ---- what the reference keeps up to iteration `last` (lane 0) ----
	int lo = 0, clo = 0, chi = 0, first_low = -1;
	double fsum = 0.0;
	for (int it = 0; it <= last; it++) {
		const bool low = bit(it);
		const float xrem = low ? s(lo) : s(n - 1 - (it - lo));
		const float xflag = low ? xrem : s(n - 1 - it);	// a high-end removal flags index size - 1 of the SHRUNK window in the original stack
		if (low) { if (first_low < 0) first_low = it; lo++; }
		if ((double) xrem >= median) chi++; else clo++;
		fsum += (double) xflag;
	}
	rej[0] += clo; rej[1] += chi;
	o.status = FAST_DONE; o.kept = n - (last + 1); o.sum = S.sum(n) - fsum;
	o.pmin = s(lo);
	o.pmax = (first_low >= 0) ? s(n - 1 - first_low) : s(n - last - 2);
}

template <int LT, int RE, int RS, int RF, int SUB, int NW>
__global__ void __launch_bounds__(NW * 32, (RE <= 24 && RE >= 20 ? 1024 : 768) / (NW * 32)) stack_lsort_kernel(const StackParams p, const __grid_constant__ CUtensorMap tmap) {
	constexpr int T = 1 << LT, Q = 32 >> LT, PXC = SUB * TILE_PX;
	extern __shared__ __align__(1024) unsigned char smem_raw[];
	const int N = p.N;
	float *rows = (float *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);	// [SUB][tile_rows][32]
	const size_t sub_floats = (size_t) ((p.tile_rows + 7) & ~7) * TILE_PX;	// sub-tiles start on 1024-byte swizzle atoms
	Sums *sums = (Sums *) (rows + SUB * sub_floats);
	int *cnt = (int *) (sums + PXC), *bad = cnt + PXC;
	unsigned long long *bar = (unsigned long long *) (bad + PXC);
	double2 *coef = (double2 *) (bar + 2);	// [N] {scale, offset | mul} when the tile holds raw samples
	const int layer = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const size_t npix = (size_t) p.H * p.W, pix0 = (size_t) blockIdx.x * PXC;

	if (p.use_tma) {
		if (threadIdx.x == 0) mbar_init(bar, 1);
		__syncthreads();
		if (threadIdx.x == 0) {
			int nsub = 0;	// sub-tiles that start inside the image (the others are never looked at)
			for (int sb = 0; sb < SUB; sb++) if (pix0 + (size_t) sb * TILE_PX < npix) nsub++;
			mbar_expect_tx(bar, (unsigned) (nsub * (size_t) p.tile_rows * TILE_PX * sizeof(float)));
			for (int sb = 0; sb < nsub; sb++)
				for (int r0 = 0; r0 < p.tile_rows; r0 += p.box_rows)
					tma_load_3d(rows + sb * sub_floats + (size_t) r0 * TILE_PX, &tmap, (int) (pix0 + sb * TILE_PX), layer, r0, bar);
		}
		if (p.normalize != SGPU_NO_NORM) {	// per-frame coefficients, staged while the tile is in flight
			const bool add = p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING;
			const double *c0 = p.scale + layer * N, *c1 = (add ? p.offset : p.mul) + layer * N;
			for (int i = threadIdx.x; i < N; i += NW * 32) coef[i] = make_double2(c0[i], c1[i]);
			__syncthreads();
		}
		mbar_wait(bar, 0);
	} else {
		for (int sb = 0; sb < SUB; sb++) {
			if (p.is_u16) load_tile_any<u16, NW>(p, rows + sb * sub_floats, layer, pix0 + sb * TILE_PX, npix, warp, lane);
			else load_tile_any<float, NW>(p, rows + sb * sub_floats, layer, pix0 + sb * TILE_PX, npix, warp, lane);
		}
		__syncthreads();
	}
	const int tma_mode = !p.use_tma ? 0 : (p.normalize == SGPU_NO_NORM ? 0 : ((p.normalize == SGPU_ADDITIVE || p.normalize == SGPU_ADDITIVE_SCALING) ? 1 : 2));

	// ---- sort: warp w works on sub-tile w % SUB, rounds w / SUB, + NW / SUB, ... ----
	{
		const int t = lane & (T - 1), g = lane >> LT;
		const unsigned gmask = (T == 32) ? 0xffffffffu : (((1u << T) - 1u) << (lane & ~(T - 1)));
		float *srows = rows + (warp % SUB) * sub_floats;	// 1024-byte aligned
		const float nullv = p.is_mean ? 0.f : __int_as_float(0x7fc00000);	// samples equal to this are dropped (nothing equals NaN)
		constexpr int NA = (T >= 8) ? 1 : 8 / T;	// distinct values of (j*T) & 7
#pragma unroll 1
		for (int r = warp / SUB; r < T; r += NW / SUB) {
			const int px = lsort_pixel<LT>(r, g);
			// byte address of (row j*T + t, px) = (la[k] + j*T*128) with k = ((j*T) & 7) / T: the 128B swizzle
			// XORs the 16-byte chunk index with row & 7 = ((j*T) & 7) | (t & 7) [T <= 8], t & 7 [T > 8]
			const char *la[NA];
			{
				const unsigned a0 = (unsigned) (t * 128) | ((unsigned) (((px >> 2) ^ t) & 7) << 4) | ((unsigned) (px & 3) << 2);
#pragma unroll
				for (int k = 0; k < NA; k++) la[k] = (const char *) srows + (a0 ^ ((unsigned) ((k * T) & 7) << 4));
			}
			float v[RE];
			int n = 0;
			float poison = 0.f;	// becomes NaN if a sample is not finite
#pragma unroll
			for (int j = 0; j < RE; j++) {
				float x = SGPU_INF;
				if (j < RS) {
					const int i = j * T + t;
					if (j < RF || i < N) {	// rows below RF are full for every stack depth this variant serves
						x = *(const float *) (la[(((j * T) & 7) / T) % NA] + j * T * 128);
						if (tma_mode) {
							const double2 c = coef[i];
							x = tma_mode == 1 ? normalize_sample<float, 1>(x, c.x, c.y) : normalize_sample<float, 2>(x, c.x, c.y);
						}
						poison = fmaf(x, 0.f, poison);
						const bool ok = x != nullv;
						n += ok;
						x = ok ? x : SGPU_INF;
					}
				}
				v[j] = x;
			}
			Sums S; S.K = 0.0; S.S1 = 0.0; S.S2 = 0.0;
			if (p.need_sums) {	// sum(x - K), sum((x - K)^2) over the valid samples, K: a sample of the group's first lane
				float Kf = __shfl_sync(gmask, v[0], 0, T);
				Kf = (Kf < SGPU_INF) ? Kf : 0.f;
				const double K = (double) Kf;
				double a1 = 0.0, a2 = 0.0, b1 = 0.0, b2 = 0.0;
#pragma unroll
				for (int j = 0; j < RS; j++) {
					const float xs = (v[j] < SGPU_INF) ? v[j] : Kf;	// dropped samples and padding count as K: y == 0
					const double y = (double) xs - K;
					if (j & 1) { b1 += y; b2 = fma(y, y, b2); } else { a1 += y; a2 = fma(y, y, a2); }
				}
				a1 += b1; a2 += b2;
#pragma unroll
				for (int m = 1; m < T; m <<= 1) { a1 += __shfl_xor_sync(gmask, a1, m); a2 += __shfl_xor_sync(gmask, a2, m); }
				S.K = K; S.S1 = a1; S.S2 = a2;
			}
			int weird = !(poison == 0.f);
#pragma unroll
			for (int m = 1; m < T; m <<= 1) { n += __shfl_xor_sync(gmask, n, m); weird |= __shfl_xor_sync(gmask, weird, m); }
			LocalNet<RS>::template run<RE>(v);
			group_merge<LT, RE>(v, t, gmask);
			__syncwarp();	// every lane of the warp has read its rows of this round's pixels
			const int pxc = (warp % SUB) * TILE_PX + px;
			if (t == 0) { cnt[pxc] = n; bad[pxc] = weird; if (p.need_sums) sums[pxc] = S; }
			// rank t*RE + j goes to row t*RE + j, whose swizzle term is (j + t*RE) & 7 = (j & 7) ^ ((t*RE) & 7):
			// RE is a multiple of 4, so t*RE & 7 is 0 or 4 and the addition is an XOR
			static_assert(RE % 4 == 0, "register slots per lane must be a multiple of 4");
			char *sa = (char *) srows + ((unsigned) (t * RE * 128) | ((unsigned) (((px >> 2) ^ ((t * RE) & 7)) & 7) << 4) | ((unsigned) (px & 3) << 2));
			const int lim = N - t * RE;
#pragma unroll
			for (int j = 0; j < RE; j++) if (j < lim) *(float *) ((char *) ((uintptr_t) sa ^ ((unsigned) (j & 7) << 4)) + j * 128) = v[j];
		}
	}
	__syncthreads();
	RejCfg rc = p.rc;
	if (p.is_mean && rc.type == SGPU_LINEARFIT) {	// table 1.f / (i + 1) of the running means, over the (now dead) coefficients
		float *xf = (float *) coef;
		for (int i = threadIdx.x; i < ((N + 7) & ~7); i += NW * 32) xf[i] = 1.f / (float) (i + 1);
		rc.xf = xf;
		__syncthreads();
	}
	// ---- GESDT: GL lanes per pixel (coop_gesdt); the pixels it does not take (too few samples, budgets that could
	// collide, very deep budgets) go through the one-thread code below, by the group's first lane ----
	constexpr int GL = 4;
	if (p.is_mean && rc.type == SGPU_GESDT && NW * 32 >= GL * PXC) {
		if ((int) threadIdx.x >= GL * PXC) return;
		unsigned *dec = (unsigned *) (coef + N);	// [SGPU_GESDT_WORDS][PXC]
		if ((int) threadIdx.x < PXC) {	// the decisions: thread i = pixel i
			const int pd = (int) threadIdx.x, nd = cnt[pd];
			int mo = (int) ((float) N * rc.s0);
			const int rem = N - nd;
			const bool ok = pix0 + pd < npix && !bad[pd] && nd >= 2 && rem < mo;
			mo -= rem;
			if (ok && !(2 * mo >= nd || nd - mo < 3) && mo >= 1 && mo <= SGPU_GESDT_COOP_MAX) {
				SubStack sd(rows + (pd >> 5) * sub_floats, pd & 31);
				if (p.is_u16) gesdt_decide<true>(sd, nd, mo, rc, sums[pd], dec + pd, PXC);
				else gesdt_decide<false>(sd, nd, mo, rc, sums[pd], dec + pd, PXC);
			}
		}
		asm volatile("bar.sync 1, %0;" :: "r"(GL * PXC) : "memory");	// (the warps past GL * PXC have left)
		const int pxi = (int) threadIdx.x / GL, q = (int) threadIdx.x % GL, sbq = pxi >> 5;
		const size_t pixq = pix0 + pxi;
		const bool valid = pixq < npix;
		SubStack st(rows + sbq * sub_floats, pxi & 31);
		const int n = cnt[pxi];
		int status = (valid && !bad[pxi]) ? FAST_DONE : FAST_FALLBACK;
		int max_out = (int) ((float) N * rc.s0);
		const int removed = N - n;
		const bool budget = status == FAST_DONE && n >= 2 && removed < max_out;
		max_out -= removed;
		const bool main_case = budget && !(2 * max_out >= n || n - max_out < 3) && max_out >= 1 && max_out <= SGPU_GESDT_COOP_MAX;
		int rejq[2] = { 0, 0 };
		FastRes o; o.status = FAST_FALLBACK; o.kept = 0; o.sum = 0.0; o.pmin = o.pmax = 0.f;
		const Sums Sq = sums[pxi];
		if (p.is_u16) coop_gesdt<true, GL>(st, n, removed, max_out, rc, Sq, q, main_case, dec + pxi, PXC, rejq, o);
		else coop_gesdt<false, GL>(st, n, removed, max_out, rc, Sq, q, main_case, dec + pxi, PXC, rejq, o);
		if (q == 0 && valid) {
			const int x = (int) (pixq % p.W);
			const double *fw = p.weights ? p.weights + (size_t) layer * N : nullptr;
			const PixWeights w = { fw, p.mask ? p.mask + pixq : nullptr, nullptr, npix, 0, p.shiftx };
			double result = 0.0;
			if (status == FAST_DONE && main_case) {
				status = (o.status == FAST_DONE && o.kept > 0) ? FAST_DONE : FAST_FALLBACK;
				if (status == FAST_DONE) {
					if (p.is_u16) { auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pixq, x); }; result = fast_combine(o, N, w, orig); }
					else { auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pixq, x); }; result = fast_combine(o, N, w, orig); }
				}
			} else if (status == FAST_DONE) {	// the cases the group does not take: the one-thread code
				const Sums *S0 = n > 0 ? &sums[pxi] : nullptr;
				if (p.is_u16) { auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pixq, x); }; result = fast_mean_and_reject<true>(st, n, N, rc, w, orig, S0, rejq, &status); }
				else { auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pixq, x); }; result = fast_mean_and_reject<false>(st, n, N, rc, w, orig, S0, rejq, &status); }
			}
			if (status == FAST_DONE) {
				if (p.is_u16) store_result<u16>(p, (size_t) layer * p.out_plane + pixq, result, rejq);
				else store_result<float>(p, (size_t) layer * p.out_plane + pixq, result, rejq);
			} else {
				rejq[0] = rejq[1] = 0;
				p.list[atomicAdd(p.list_count, 1u)] = ((unsigned) layer << 30) | (unsigned) pixq;
			}
		} else { rejq[0] = rejq[1] = 0; }
		__syncwarp();
		unsigned lo = rejq[0], hi = rejq[1];
		lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
		if (lane == 0 && (lo | hi)) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
		}
		return;
	}
	if (threadIdx.x >= PXC) return;

	// ---- reject / combine: thread i of the CTA owns pixel i of the tile ----
	const int sb = threadIdx.x >> 5;
	const size_t pix = pix0 + threadIdx.x;
	int rej[2] = { 0, 0 };
	if (pix < npix) {
		SubStack st(rows + sb * sub_floats, lane);
		const int n = cnt[threadIdx.x];
		int status = bad[threadIdx.x] ? FAST_FALLBACK : FAST_DONE;
		double result = 0.0;
		if (status == FAST_DONE) {
			if (p.is_mean) {
				const int x = (int) (pix % p.W);
				const double *fw = p.weights ? p.weights + (size_t) layer * N : nullptr;
				// frame weights, and the pixel's feather-mask samples where the block carries masks (median_and_mean.c:1060-1066)
				const PixWeights w = { fw, p.mask ? p.mask + pix : nullptr, nullptr, npix, 0, p.shiftx };
				const Sums *S0 = (p.need_sums && n > 0) ? &sums[threadIdx.x] : nullptr;
				if (p.is_u16) {
					auto orig = [&](int ff) { return (float) load_norm<u16>(p, ff, layer, pix, x); };
					result = fast_mean_and_reject<true>(st, n, N, rc, w, orig, S0, rej, &status);
				} else {
					auto orig = [&](int ff) { return load_norm<float>(p, ff, layer, pix, x); };
					result = fast_mean_and_reject<false>(st, n, N, rc, w, orig, S0, rej, &status);
				}
			} else {	// all N samples, zeros included (median_and_mean.c:1711-1715)
				result = p.is_u16 ? median_sorted<true>(st, 0, N) : median_sorted<false>(st, 0, N);
			}
		}
		if (status == FAST_DONE) {
			if (p.is_u16) store_result<u16>(p, (size_t) layer * p.out_plane + pix, result, rej);
			else store_result<float>(p, (size_t) layer * p.out_plane + pix, result, rej);
		} else {
			rej[0] = rej[1] = 0;
			p.list[atomicAdd(p.list_count, 1u)] = ((unsigned) layer << 30) | (unsigned) pix;
		}
	}
	if (p.is_mean && p.rc.type != SGPU_NO_REJEC) {
		__syncwarp();
		unsigned lo = rej[0], hi = rej[1];
		lo = __reduce_add_sync(0xffffffffu, lo); hi = __reduce_add_sync(0xffffffffu, hi);
		if (lane == 0 && (lo | hi)) {
			atomicAdd(&p.irej[layer * 2], (unsigned long long) lo);
			atomicAdd(&p.irej[layer * 2 + 1], (unsigned long long) hi);
		}
	}
}

}  // namespace sgpu

template <int LT, int RE, int RS, int RF, int SUB, int NW>
int sgpu_launch_lsort_variant(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm);

#define SGPU_LSORT_VARIANT(LT, RE, RS, RF, SUB, NW) \
	template <> int sgpu_launch_lsort_variant<LT, RE, RS, RF, SUB, NW>(sgpu_ctx *ctx, const sgpu::StackParams &p, const CUtensorMap *tmap, dim3 grid, size_t shm) { \
		auto kern = sgpu::stack_lsort_kernel<LT, RE, RS, RF, SUB, NW>; \
		const CUtensorMap tm = tmap ? *tmap : CUtensorMap{}; \
		if (shm > 48 * 1024) SGPU_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) shm)); \
		kern<<<grid, NW * 32, shm, ctx->stream>>>(p, tm); \
		ctx->launches++; \
		SGPU_CUDA(ctx, cudaGetLastError()); \
		return SGPU_OK; \
	}
