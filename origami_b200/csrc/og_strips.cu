// Strip extraction with the binarizer fused in (A8 + A9 + A10 + A12): the engine behind
// og_extract_lines.
//
// Reference: Line.image = Line.dewarped_image + binarizer(im), origami/core/block.py:129-143 and
// :250-259; LineExtractor.__call__ maps it over the lines of a page with a thread pool,
// origami/batch/core/lines.py:131-149.
//
// One CTA owns a text line (lines are handed out through a ticket counter).  Once per line it
// plans every 64-column chunk (the rectangle of dense-map nodes under it, an error radius for the
// screening below, whether any sample can leave the map or the page) and tabulates the strip rows
// and the y axis of cv2.resize for the dense rows under the line.  Then, per chunk:
//   1. the chunk's rectangle of dense-map nodes is recomputed from the sparse grid into shared
//      memory as float32 (x, y) pairs, with the exact cv2.resize arithmetic -- no dense map is
//      read (SURVEY.md section 8(d) budgets zero bytes for it);
//   2. every lattice point of the chunk is located in that rectangle with 32-bit fixed-point
//      coordinates relative to the rectangle's origin (9.23 bits: no float64 add, no conversion
//      in the sample loop); its four nodes are four 8-byte shared-memory loads;
//   3. what cv2.remap consumes of SciPy's float64 blend X is cvRound(32 * fl32(X)).  As in the
//      screened kernel of og_lines.cu it is decided from a float32 interval that provably contains
//      X, evaluated here on (x, y) pairs with Blackwell's packed FADD2 / FMUL2 / FFMA2; the
//      error radius uses one bound G on the node differences for the whole chunk, taken from the
//      sparse cells under it.  Samples the interval cannot decide, or that leave the interior of
//      the map or of the page, take the float64 sequence of the reference on the same nodes;
//   4. the four source bytes are gathered and blended with OpenCV's integer weights, the gray
//      byte goes to the strip (row 0 = top of the text) and, for Otsu, into a replicated
//      shared-memory histogram.
// After the last chunk the CTA computes the line's Otsu threshold from the histogram and
// binarizes its own strip in place (the bytes are still in L2), so a binarized strip costs no
// second kernel and no second pass over HBM.
#include "og_common.cuh"
#include "og_internal.cuh"

namespace {

constexpr int FT = 256;           // threads per CTA
constexpr int FWARPS = FT / 32;
constexpr int FCW = 64;           // strip columns per chunk
constexpr int COLG = FCW / 32;    // column groups of 32 lanes
constexpr int ROWG = FWARPS / COLG;  // row groups
#ifndef OG_F_NODE_CAP
#define OG_F_NODE_CAP 5120
#endif
constexpr int NODE_CAP = OG_F_NODE_CAP;  // float2 nodes per staged rectangle
constexpr int NR_CAP = 160;       // dense rows under a row band of a line
constexpr int ROW_CAP = 128;      // strip rows per band
constexpr int PLAN_CAP = 32;      // chunks planned at a time
#ifndef OG_F_HREP
#define OG_F_HREP 4
#endif
constexpr int FHREP = OG_F_HREP;  // histogram replicas ([bin][lane % FHREP])
#ifndef OG_F_NB
#define OG_F_NB 4
#endif
constexpr int FNB = OG_F_NB;      // strip rows a lane samples together
#ifndef OG_F_MINB
#define OG_F_MINB 3
#endif

typedef unsigned long long u64;

__device__ __forceinline__ u64 pk(float lo, float hi) {
	u64 r;
	asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
	return r;
}
__device__ __forceinline__ float lo_f(u64 v) { return __uint_as_float((unsigned)v); }
__device__ __forceinline__ float hi_f(u64 v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ u64 add2(u64 a, u64 b) {
	u64 r;
	asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ u64 sub2(u64 a, u64 b) {
	u64 r;
	asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ u64 mul2(u64 a, u64 b) {
	u64 r;
	asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
	return r;
}
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) {
	u64 r;
	asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
	return r;
}
// fl(fl(a * wa) + fl(b * wb)) per component -- cv2.resize's HResizeLinear / VResizeLinear step on an
// (x, y) pair.  The two sums are scalar adds on purpose: ptxas contracts mul.rn.f32x2 followed by
// add.rn.f32x2 into one FFMA2 (seen in the SASS, --fmad=false notwithstanding), which skips the
// rounding of the second product and moves every fourth node by an ulp.
__device__ __forceinline__ u64 lerp2_cv(u64 a, u64 wa, u64 b, u64 wb) {
	const u64 pa = mul2(a, wa), pb = mul2(b, wb);
	return pk(__fadd_rn(lo_f(pa), lo_f(pb)), __fadd_rn(hi_f(pa), hi_f(pb)));
}

struct FusedParams {
	const uint8_t *src;
	int H, W, pitch;
	const float2 *grid;
	int gh, gw, cell;
	const og_line *lines;
	int n;
	const int32_t *steps;
	const int64_t *pix_offsets;
	int *ticket;
	uint8_t *out;   // the strips the caller asked for (binarized, or gray for OG_BIN_NONE)
	uint8_t *gray;  // where the gray bytes are written first (== out unless the caller wants both)
	int64_t capacity;
	int32_t *thresholds;
	int W1, H1, Hd, Wd;
	double inv_cell;
	float g_scale;  // 1.01 / cell
};

// Chunk flags
constexpr int RF_STAGED = 1;    // the rectangle fits: samples are located in shared memory
constexpr int RF_INSIDE = 2;    // no clamping at the border of the map: every cell index is valid by construction
constexpr int RF_TAPS_IN = 4;   // every tap of every sample lies inside the page

struct ChunkPlan {
	int R0, C0, NR, NC;  // rectangle of dense-map nodes: rows R0 .. R0 + NR - 1, columns C0 .. C0 + NC - 1
	int NCP;             // row pitch of the staged rectangle in nodes (NC rounded up to even)
	int flags;
	float gbound;        // bound on the difference of neighbouring nodes in the rectangle
	int yo;              // R0 - first dense row of the band's y-axis table
	int groups, band;    // staging: row groups and rows per group
	float inv_pairs;     // 1 / (NCP / 2)
	int pad_;
};

struct __align__(16) FusedSmem {
	float2 nodes[NODE_CAP];
	float4 ytab_b[NR_CAP];        // (b0, b0, b1, b1) of cv2.resize's vertical pass per dense row under the band
	int2 ytab_r[NR_CAP + 1];      // its two sparse rows (clamped); one sentinel entry behind the last row
	int2 rowfx[ROW_CAP + 16];     // (ys[r] + p0) of the band's strip rows relative to the band's origin, 9.23 fixed point
	ChunkPlan plans[PLAN_CAP];
	unsigned int hist[256 * FHREP];
	unsigned int hsum[256];
	double lin[4];                // the four np.linspace steps of the line
	int band[4];                  // first dense row and number of dense rows under the band
	int line;
	int thr;
};

// One sample in the reference's arithmetic only: float64 lattice point, SciPy's float64 blend on
// the exact float32 nodes (from the staged rectangle when it holds them, else recomputed from the
// sparse grid), float32 cast, cv2.remap's fixed point, per-tap border test.
__device__ __noinline__ int sample_exact(const FusedParams &p, const FusedSmem &s, const ChunkPlan &cp, bool staged,
	const og_lin &lx, const og_lin &ly, double cpx, double cpy, int r, int cc) {
	const og_d2 yr = og_linspace_at(ly, r);
	const og_d2 xc = og_linspace_at(lx, cc);
	const double px = __dadd_rn(__dadd_rn(yr.x, cpx), xc.x), py = __dadd_rn(__dadd_rn(yr.y, cpy), xc.y);
	double fy = floor(py), fx = floor(px);
	fy = fmin(fmax(fy, 0.0), (double)(p.Hd - 2));
	fx = fmin(fmax(fx, 0.0), (double)(p.Wd - 2));
	if (!(fy == fy) || !(fx == fx)) fy = fx = 0.0;  // NaN coordinates: any cell, the blend stays NaN
	const int i0 = (int)fy, i1 = (int)fx;
	const double a = __dsub_rn(py, fy), b = __dsub_rn(px, fx);
	float2 v00, v01, v10, v11;
	const int ri = i0 - cp.R0, ci = i1 - cp.C0;
	if (staged && ri >= 0 && ci >= 0 && ri <= cp.NR - 2 && ci <= cp.NC - 2) {
		const float2 *nd = s.nodes + ri * cp.NCP + ci;
		v00 = nd[0]; v01 = nd[1]; v10 = nd[cp.NCP]; v11 = nd[cp.NCP + 1];
	} else {
		v00 = og_dense_node(p.grid, p.gh, p.gw, p.inv_cell, i0, i1);
		v01 = og_dense_node(p.grid, p.gh, p.gw, p.inv_cell, i0, i1 + 1);
		v10 = og_dense_node(p.grid, p.gh, p.gw, p.inv_cell, i0 + 1, i1);
		v11 = og_dense_node(p.grid, p.gh, p.gw, p.inv_cell, i0 + 1, i1 + 1);
	}
	const og_d2 w = og_blend_f64(v00, v01, v10, v11, a, b);
	const float gx = __double2float_rn(w.x), gy = __double2float_rn(w.y);
	if (!(gx == gx) || !(gy == gy)) return 0;  // cvRound(NaN) is INT_MIN on the reference's hosts: every tap is outside
	const int sx[1] = {__float2int_rn(__fmul_rn(gx, 32.f))}, sy[1] = {__float2int_rn(__fmul_rn(gy, 32.f))};
	og_taps t[1];
	og_load_taps_fixed<1>(p.src, p.H, p.W, p.pitch, sx, sy, t);
	return og_blend_taps(t[0]);
}

#define OG_RINT_MAGIC_F 12582912.0f
#define OG_RINT_BITS_I 0x4B400000

// The rectangle of dense-map nodes under strip rows [ra, rb) x columns [c0, c0 + ncols) of a line.
// The lattice is affine and np.linspace is monotone, so its extremes are at the four corners; one
// node of margin on every side absorbs the rounding of the fixed-point coordinates.
__device__ __forceinline__ void plan_rect(const FusedParams &p, const og_lin &lx, const og_lin &ly, double cpx, double cpy,
	int ra, int rb, int c0, int ncols, ChunkPlan &cp) {
	const og_d2 xa = og_linspace_at(lx, c0), xb = og_linspace_at(lx, c0 + ncols - 1);
	const og_d2 ya0 = og_linspace_at(ly, ra), yb0 = og_linspace_at(ly, rb - 1);
	const double yax = __dadd_rn(ya0.x, cpx), yay = __dadd_rn(ya0.y, cpy), ybx = __dadd_rn(yb0.x, cpx), yby = __dadd_rn(yb0.y, cpy);
	const double x0 = __dadd_rn(yax, xa.x), x1 = __dadd_rn(yax, xb.x), x2 = __dadd_rn(ybx, xa.x), x3 = __dadd_rn(ybx, xb.x);
	const double y0 = __dadd_rn(yay, xa.y), y1 = __dadd_rn(yay, xb.y), y2 = __dadd_rn(yby, xa.y), y3 = __dadd_rn(yby, xb.y);
	const double xmin = fmin(fmin(x0, x1), fmin(x2, x3)), xmax = fmax(fmax(x0, x1), fmax(x2, x3));
	const double ymin = fmin(fmin(y0, y1), fmin(y2, y3)), ymax = fmax(fmax(y0, y1), fmax(y2, y3));
	const int r0u = (int)floor(ymin) - 1, r1u = (int)floor(ymax) + 2;
	const int c0u = (int)floor(xmin) - 1, c1u = (int)floor(xmax) + 2;
	cp.R0 = min(max(r0u, 0), p.Hd - 2);
	const int R1 = min(max(r1u, 1), p.Hd - 1);
	cp.C0 = min(max(c0u, 0), p.Wd - 2);
	const int C1 = min(max(c1u, 1), p.Wd - 1);
	cp.NR = max(R1 - cp.R0 + 1, 2);
	cp.NC = max(C1 - cp.C0 + 1, 2);
	cp.flags = (r0u >= 0 && c0u >= 0 && r1u <= p.Hd - 1 && c1u <= p.Wd - 1) ? RF_INSIDE : 0;
	cp.gbound = 0.f;
	cp.NCP = (cp.NC + 1) & ~1;
	const int pairs = cp.NCP >> 1;
	cp.groups = max(1, min(FT / pairs, cp.NR));
	cp.band = (cp.NR + cp.groups - 1) / cp.groups;
	cp.inv_pairs = 1.f / (float)pairs;
	cp.yo = 0;
}

// The batch of FNB samples again, for a warp in which some lane's sample was not decided by the
// screening: decided samples fetch their taps, the others take the float64 path.  Returns the FNB
// gray values packed into one word.  Kept out of line so that the common path stays small.
__device__ __noinline__ unsigned int sample_batch_slow(const FusedParams &p, const FusedSmem &s, const ChunkPlan &cp,
	const og_lin &lx, const og_lin &ly, double cpx, double cpy, int r, int r_last, int cc,
	int sx0, int sy0, int sx1, int sy1, int sx2, int sy2, int sx3, int sy3, unsigned okmask) {
	const int sx[4] = {sx0, sx1, sx2, sx3}, sy[4] = {sy0, sy1, sy2, sy3};
	unsigned int packed = 0;
#pragma unroll
	for (int q = 0; q < 4; ++q) {
		int v;
		if (okmask & (1u << q)) {
			const uint8_t *tp = p.src + ((sy[q] >> 5) * p.pitch + (sx[q] >> 5));
			og_taps t;
			t.fx = sx[q] & 31;
			t.fy = sy[q] & 31;
			t.p00 = __ldg(tp);
			t.p01 = __ldg(tp + 1);
			t.p10 = __ldg(tp + p.pitch);
			t.p11 = __ldg(tp + p.pitch + 1);
			v = og_blend_taps(t);
		} else {
			v = sample_exact(p, s, cp, true, lx, ly, cpx, cpy, min(r + q, r_last), cc);
		}
		packed |= (unsigned int)v << (8 * q);
	}
	return packed;
}

// Samples of one warp's share of a chunk: 32 columns x rows [r_beg, r_end) of the band.
// CHECKED = false: the plan guarantees that every cell index is valid and every tap inside the page.
template <int BIN, bool CHECKED>
__device__ __forceinline__ void sample_rows(const FusedParams &p, FusedSmem &s, const ChunkPlan &cp,
	const og_lin &lx, const og_lin &ly, double cpx, double cpy,
	int ra, int r_beg, int r_end, int cc_clamped, bool col_ok, int XCx, int XCy,
	uint8_t *outp, int xsteps, unsigned int *myhist) {
	static_assert(FNB == 4, "sample_batch_slow takes four samples");
	const int NCP = cp.NCP, NRm = cp.NR - 2, NCm = cp.NC - 2;
	const float eps = __fmaf_rn(cp.gbound, 0x1p-19f, 0x1p-30f);  // 32 u G + 2^-30
	const u64 epsp = pk(eps, eps), c32 = pk(32.f, 32.f), magic = pk(OG_RINT_MAGIC_F, OG_RINT_MAGIC_F);
	const uint8_t *__restrict__ src = p.src;
	const int spitch = p.pitch;
	const int2 *rfp = s.rowfx + (r_beg - ra);
	const unsigned char *nbase = reinterpret_cast<const unsigned char *>(s.nodes);
	const int rowbytes = NCP * 8;
	for (int r = r_beg; r < r_end; r += FNB, rfp += FNB) {
		int sx[FNB], sy[FNB];
		unsigned okmask = 0;
#pragma unroll
		for (int q = 0; q < FNB; ++q) {
			const int2 rf = rfp[q];  // the table is padded with copies of the last row
			const int fxx = rf.x + XCx, fyy = rf.y + XCy;
			const int i1 = fxx >> 23, i0 = fyy >> 23;
			bool ok = true;
			int off = i0 * rowbytes + i1 * 8;
			if (CHECKED) {
				ok = (unsigned)i0 <= (unsigned)NRm && (unsigned)i1 <= (unsigned)NCm;
				off = ok ? off : 0;
			}
			const float b = __fsub_rn(__int_as_float((fxx & 0x7FFFFF) | 0x3F800000), 1.f);
			const float a = __fsub_rn(__int_as_float((fyy & 0x7FFFFF) | 0x3F800000), 1.f);
			const u64 *n0 = reinterpret_cast<const u64 *>(nbase + off);
			const u64 *n1 = reinterpret_cast<const u64 *>(nbase + off + rowbytes);
			const u64 v00 = n0[0], v01 = n0[1], v10 = n1[0], v11 = n1[1];
			const u64 d01 = sub2(v01, v00), d10 = sub2(v10, v00), d11 = sub2(v11, v10);
			const u64 ap = pk(a, a), bp = pk(b, b);
			const u64 del = fma2(ap, fma2(bp, sub2(d11, d01), d10), mul2(bp, d01));
			const u64 ul = fma2(add2(v00, sub2(del, epsp)), c32, magic);
			const u64 uh = fma2(add2(v00, add2(del, epsp)), c32, magic);
			sx[q] = (int)(unsigned)ul - OG_RINT_BITS_I;
			sy[q] = (int)(unsigned)(ul >> 32) - OG_RINT_BITS_I;
			// float compares: a NaN is never decided.  The mantissa holds cvRound(32 x) while |32 x| < 2^22,
			// which the page test (or the plan's source box) implies.
			ok = ok && lo_f(ul) == lo_f(uh) && hi_f(ul) == hi_f(uh);
			if (CHECKED) ok = ok && (unsigned)(sx[q] >> 5) < (unsigned)p.W1 && (unsigned)(sy[q] >> 5) < (unsigned)p.H1;
			okmask |= ok ? (1u << q) : 0u;
		}
		unsigned int packed;
		if (__all_sync(0xffffffffu, okmask == (1u << FNB) - 1)) {
			og_taps t[FNB];
#pragma unroll
			for (int q = 0; q < FNB; ++q) {
				const uint8_t *tp = src + ((sy[q] >> 5) * spitch + (sx[q] >> 5));
				t[q].fx = sx[q] & 31;
				t[q].fy = sy[q] & 31;
				t[q].p00 = __ldg(tp);
				t[q].p01 = __ldg(tp + 1);
				t[q].p10 = __ldg(tp + spitch);
				t[q].p11 = __ldg(tp + spitch + 1);
			}
			packed = 0;
#pragma unroll
			for (int q = 0; q < FNB; ++q) packed |= (unsigned int)og_blend_taps(t[q]) << (8 * q);
		} else {
			packed = sample_batch_slow(p, s, cp, lx, ly, cpx, cpy, r, r_end - 1, cc_clamped,
				sx[0], sy[0], sx[1], sy[1], sx[2], sy[2], sx[3], sy[3], okmask);
		}
#pragma unroll
		for (int q = 0; q < FNB; ++q) {
			const unsigned int v = (packed >> (8 * q)) & 255u;
			if (col_ok && r + q < r_end) {
				*outp = (uint8_t)v;
				if (BIN == OG_BIN_OTSU) atomicAdd(myhist + v * FHREP, 1u);
			}
			outp -= xsteps;
		}
	}
}

template <int BIN>
__global__ void __launch_bounds__(FT, OG_F_MINB) og_strips_fused_kernel(const __grid_constant__ FusedParams p) {
	extern __shared__ __align__(16) unsigned char smem_raw[];
	FusedSmem &s = *reinterpret_cast<FusedSmem *>(smem_raw);
	const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
	const int cg = wid % COLG, rg = wid / COLG;
	unsigned int *myhist = s.hist + (lane & (FHREP - 1));

	if (BIN == OG_BIN_OTSU)
		for (int k = tid; k < 256 * FHREP; k += FT) s.hist[k] = 0;

	for (;;) {
		__syncthreads();  // the previous line is finished with the shared tables
		if (tid == 0) s.line = atomicAdd(p.ticket, 1);
		__syncthreads();
		const int li = s.line;
		if (li >= p.n) {
			// every CTA draws exactly one ticket past the end; the last one leaves the counter at zero
			if (tid == 0 && li == p.n + (int)gridDim.x - 1) *p.ticket = 0;
			break;
		}
		const int xsteps = __ldg(p.steps + 2 * li), ysteps = __ldg(p.steps + 2 * li + 1);
		const long long pix0 = __ldg(p.pix_offsets + li);
		const long long npix = (long long)xsteps * ysteps;
		if (xsteps <= 0 || ysteps <= 0 || pix0 + npix > p.capacity) continue;  // a strip that does not fit is skipped whole
		const og_line *ln = p.lines + li;
		const double cpx = ln->cp[0], cpy = ln->cp[1];
		const double crx = ln->cright[0], cry = ln->cright[1], upx = ln->up[0], upy = ln->up[1];
		if (tid < 4) {  // the four float64 divisions of the two np.linspace calls, once per line
			const double stop = tid == 0 ? crx : tid == 1 ? cry : tid == 2 ? upx : upy;
			const int num = tid < 2 ? xsteps : ysteps;
			s.lin[tid] = num > 1 ? __ddiv_rn(stop, (double)(num - 1)) : stop;
		}
		__syncthreads();
		og_lin lx, ly;
		lx.num = xsteps; lx.stopx = crx; lx.stopy = cry; lx.div = (double)(xsteps - 1);
		lx.stepx = s.lin[0]; lx.stepy = s.lin[1];
		lx.any_zero = xsteps > 1 && (lx.stepx == 0.0 || lx.stepy == 0.0);
		ly.num = ysteps; ly.stopx = upx; ly.stopy = upy; ly.div = (double)(ysteps - 1);
		ly.stepx = s.lin[2]; ly.stepy = s.lin[3];
		ly.any_zero = ysteps > 1 && (ly.stepx == 0.0 || ly.stepy == 0.0);
		// all samples are cp + s * cright + t * up with s, t in [0, 1]; beyond this bound (NaN and
		// inf included) nothing is staged and every sample takes the float64 path
		const bool sane = fabs(cpx) + fabs(cpy) + fabs(crx) + fabs(cry) + fabs(upx) + fabs(upy) < 1.0e6;
		uint8_t *gline = p.gray + pix0;
		const int nchunks = (xsteps + FCW - 1) / FCW;

		for (int ra = 0; ra < ysteps; ra += ROW_CAP) {  // one band for any OCR strip height
			const int rb = min(ysteps, ra + ROW_CAP), nrows = rb - ra;
			const int rpg = (nrows + ROWG - 1) / ROWG;
			const int r_beg = ra + rg * rpg, r_end = min(rb, r_beg + rpg);
			const int nfx = ROWG * ((rpg + FNB - 1) / FNB * FNB);
			// origin of the band's fixed-point coordinates: the integer part of its first lattice row
			const og_d2 y0 = og_linspace_at(ly, ra);
			const double LX = floor(__dadd_rn(y0.x, cpx)), LY = floor(__dadd_rn(y0.y, cpy));
			for (int kbase = 0; kbase < nchunks; kbase += PLAN_CAP) {
				const int nplan = min(PLAN_CAP, nchunks - kbase);
				// ---- plan, once per band: the strip rows in fixed point, the rectangle of every chunk, the
				// dense rows under the whole band
				__syncthreads();
				bool far = false;
				if (kbase == 0)
					for (int r = tid; r < nfx; r += FT) {
						const og_d2 yr = og_linspace_at(ly, ra + min(r, nrows - 1));  // (ys + p0), block.py:241; padded with the last row
						const double qx = __dsub_rn(__dadd_rn(yr.x, cpx), LX), qy = __dsub_rn(__dadd_rn(yr.y, cpy), LY);
						far = far || !(fabs(qx) < 120.0 && fabs(qy) < 120.0);
						s.rowfx[r] = make_int2(__double2int_rn(__dmul_rn(qx, 8388608.0)), __double2int_rn(__dmul_rn(qy, 8388608.0)));
					}
				if (tid >= 128 && tid < 128 + nplan) {
					const int k = tid - 128, c0 = (kbase + k) * FCW;
					ChunkPlan cp;
					cp.R0 = cp.C0 = 0; cp.NR = cp.NC = cp.NCP = 2; cp.flags = 0; cp.gbound = 0.f;
					cp.groups = cp.band = 1; cp.inv_pairs = 1.f; cp.yo = 0;
					if (sane) plan_rect(p, lx, ly, cpx, cpy, ra, rb, c0, min(FCW, xsteps - c0), cp);
					s.plans[k] = cp;
				}
				if (tid == 255) {
					ChunkPlan cp;
					cp.R0 = 0; cp.NR = NR_CAP + 1;
					if (sane) plan_rect(p, lx, ly, cpx, cpy, ra, rb, kbase * FCW, min(nplan * FCW, xsteps - kbase * FCW), cp);
					s.band[0] = cp.R0;
					s.band[1] = cp.NR;
				}
				far = __syncthreads_or(far) != 0;
				if (kbase == 0 && tid == 0) s.band[2] = far;
				const int RL0 = s.band[0], NRL = s.band[1];
				if (sane && NRL <= NR_CAP) {
					for (int i = tid; i <= NRL; i += FT) {
						if (i == NRL) {
							s.ytab_r[i] = make_int2(-2, -2);  // sentinel: ends the last run of rows
							break;
						}
						const og_axis ay = og_resize_axis(RL0 + i, p.inv_cell);
						const float b1 = ay.f, b0 = __fsub_rn(1.f, ay.f);
						s.ytab_b[i] = make_float4(b0, b0, b1, b1);
						s.ytab_r[i] = make_int2(min(max(ay.i, 0), p.gh - 1), min(max(ay.i + 1, 0), p.gh - 1));
					}
					// Bound on the node differences (G) and on the node values (source box) of every chunk
					// from the sparse cells under its rectangle, one warp per chunk: dense nodes are convex
					// combinations of sparse nodes up to a few float32 roundings, and neighbouring dense
					// nodes differ by at most a sparse difference divided by the cell size (plus those
					// roundings).
					for (int k = wid; k < nplan; k += FWARPS) {
						const ChunkPlan cp = s.plans[k];
						int flags = cp.flags;
						float gbound = 0.f;
						if (cp.NR * cp.NCP <= NODE_CAP && cp.NC <= 130) {
							flags |= RF_STAGED;
							const og_axis ac0 = og_resize_axis_x(cp.C0, p.inv_cell, p.gw), ac1 = og_resize_axis_x(cp.C0 + cp.NC - 1, p.inv_cell, p.gw);
							const og_axis ar0 = og_resize_axis(cp.R0, p.inv_cell), ar1 = og_resize_axis(cp.R0 + cp.NR - 1, p.inv_cell);
							const int SC0 = ac0.i, NSC = min(ac1.i + 1, p.gw - 1) - SC0 + 1;
							const int SR0 = min(max(ar0.i, 0), p.gh - 1), NSR = min(max(ar1.i + 1, 0), p.gh - 1) - SR0 + 1;
							float dmax = 0.f, vx0 = 3.0e38f, vx1 = -3.0e38f, vy0 = 3.0e38f, vy1 = -3.0e38f;
							bool finite = true;
							for (int e = lane; e < NSR * NSC; e += 32) {
								const int r = e / NSC, c = e - r * NSC;
								const float2 v = __ldg(p.grid + (SR0 + r) * p.gw + SC0 + c);
								vx0 = fminf(vx0, v.x); vx1 = fmaxf(vx1, v.x);
								vy0 = fminf(vy0, v.y); vy1 = fmaxf(vy1, v.y);
								finite = finite && fabsf(v.x) < 3.0e38f && fabsf(v.y) < 3.0e38f;  // false for NaN and inf
								if (c + 1 < NSC) {
									const float2 w = __ldg(p.grid + (SR0 + r) * p.gw + SC0 + c + 1);
									dmax = fmaxf(dmax, fmaxf(fabsf(__fsub_rn(w.x, v.x)), fabsf(__fsub_rn(w.y, v.y))));
								}
								if (r + 1 < NSR) {
									const float2 w = __ldg(p.grid + (SR0 + r + 1) * p.gw + SC0 + c);
									dmax = fmaxf(dmax, fmaxf(fabsf(__fsub_rn(w.x, v.x)), fabsf(__fsub_rn(w.y, v.y))));
								}
							}
#pragma unroll
							for (int o = 16; o > 0; o >>= 1) {
								dmax = fmaxf(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
								vx0 = fminf(vx0, __shfl_xor_sync(0xffffffffu, vx0, o));
								vx1 = fmaxf(vx1, __shfl_xor_sync(0xffffffffu, vx1, o));
								vy0 = fminf(vy0, __shfl_xor_sync(0xffffffffu, vy0, o));
								vy1 = fmaxf(vy1, __shfl_xor_sync(0xffffffffu, vy1, o));
							}
							finite = __all_sync(0xffffffffu, finite);
							// non-finite nodes: an infinite radius, so that no sample is decided in float32
							gbound = finite ? __fmaf_rn(dmax, p.g_scale, 0.03125f) : __int_as_float(0x7f800000);
							if (finite && (flags & RF_INSIDE) && vx0 >= 1.f && vy0 >= 1.f && vx1 <= (float)(p.W - 3) && vy1 <= (float)(p.H - 3))
								flags |= RF_TAPS_IN;
						}
						if (lane == 0) {
							s.plans[k].flags = flags;
							s.plans[k].gbound = gbound;
							s.plans[k].yo = cp.R0 - RL0;
						}
					}
				}
				// (the first chunk's barrier below publishes the tables)

				for (int k = 0; k < nplan; ++k) {
					const int c0 = (kbase + k) * FCW, ncols = min(FCW, xsteps - c0);
					__syncthreads();  // plans / tables written; the previous chunk's samples are done with the nodes
					const ChunkPlan cp = s.plans[k];
					const bool staged = sane && NRL <= NR_CAP && s.band[2] == 0 && (cp.flags & RF_STAGED) != 0;
					if (staged) {
						// ---- stage 1: the chunk's rectangle of dense-map nodes.  A thread owns a pair of
						// columns and a band of rows: HResizeLinear of the two sparse rows above and below
						// (kept in registers while the rows stay in the same sparse cell), then
						// VResizeLinear per row, written as one 16-byte store.
						const int pairs = cp.NCP >> 1;
						const int g = __float2int_rz(__fmul_rn((float)tid + 0.5f, cp.inv_pairs));
						const int jp = tid - g * pairs;
						if (g < cp.groups) {
							const int ibeg = g * cp.band, iend = min(cp.NR, ibeg + cp.band);
							const int ja = 2 * jp, jb = min(ja + 1, cp.NC - 1);  // an odd rectangle repeats its last column
							const og_axis axa = og_resize_axis_x(cp.C0 + ja, p.inv_cell, p.gw), axb = og_resize_axis_x(cp.C0 + jb, p.inv_cell, p.gw);
							const int ca1 = min(axa.i + 1, p.gw - 1), cb1 = min(axb.i + 1, p.gw - 1);
							const u64 wa0 = pk(__fsub_rn(1.f, axa.f), __fsub_rn(1.f, axa.f)), wa1 = pk(axa.f, axa.f);
							const u64 wb0 = pk(__fsub_rn(1.f, axb.f), __fsub_rn(1.f, axb.f)), wb1 = pk(axb.f, axb.f);
							const u64 *g64 = reinterpret_cast<const u64 *>(p.grid);
							ulonglong2 *nd = reinterpret_cast<ulonglong2 *>(s.nodes + ibeg * cp.NCP + ja);
							const int ndstep = cp.NCP >> 1;  // in 16-byte units
							const int2 *yr = s.ytab_r + cp.yo;
							const float4 *yb = s.ytab_b + cp.yo;
							int i = ibeg;
							int pr1 = -1;
							u64 ua = 0, va = 0, ub = 0, vb = 0;
							while (i < iend) {
								const int2 rr = yr[i];
								if (rr.x == pr1) {
									ua = va; ub = vb;
								} else {
									const u64 *row = g64 + rr.x * p.gw;
									ua = lerp2_cv(__ldg(row + axa.i), wa0, __ldg(row + ca1), wa1);
									ub = lerp2_cv(__ldg(row + axb.i), wb0, __ldg(row + cb1), wb1);
								}
								if (rr.y == rr.x) {
									va = ua; vb = ub;
								} else {
									const u64 *row = g64 + rr.y * p.gw;
									va = lerp2_cv(__ldg(row + axa.i), wa0, __ldg(row + ca1), wa1);
									vb = lerp2_cv(__ldg(row + axb.i), wb0, __ldg(row + cb1), wb1);
								}
								pr1 = rr.y;
								do {  // the rows of this sparse cell
									const float4 bw = yb[i];
									const u64 w0 = pk(bw.x, bw.y), w1 = pk(bw.z, bw.w);
									ulonglong2 o;
									o.x = lerp2_cv(ua, w0, va, w1);
									o.y = lerp2_cv(ub, w0, vb, w1);
									*nd = o;
									nd += ndstep;
									++i;
								} while (i < iend && yr[i].x == rr.x);
							}
						}
					}
					__syncthreads();

					// ---- stage 2: samples.  Warp = (column group, row group); a lane walks its column.
					const int cc = c0 + cg * 32 + lane;
					const bool col_ok = cg * 32 + lane < ncols;
					if (r_beg < r_end && cg * 32 < ncols) {
						uint8_t *outp = gline + (long long)(ysteps - 1 - r_beg) * xsteps + cc;  // pixels[::-1, :]   (block.py:257)
						const int ccc = min(cc, xsteps - 1);
						if (!staged) {
							if (col_ok)
								for (int r = r_beg; r < r_end; ++r) {
									const int v = sample_exact(p, s, cp, false, lx, ly, cpx, cpy, r, cc);
									*outp = (uint8_t)v;
									outp -= xsteps;
									if (BIN == OG_BIN_OTSU) atomicAdd(myhist + v * FHREP, 1u);
								}
						} else {
							// this lane's column relative to the rectangle, minus the band origin the row table
							// is relative to: (row term) + (column term) = lattice point - rectangle origin
							const og_d2 xcc = og_linspace_at(lx, ccc);
							const int XCx = __double2int_rn(__dmul_rn(__dadd_rn(xcc.x, __dsub_rn(LX, (double)cp.C0)), 8388608.0));
							const int XCy = __double2int_rn(__dmul_rn(__dadd_rn(xcc.y, __dsub_rn(LY, (double)cp.R0)), 8388608.0));
							if ((cp.flags & (RF_INSIDE | RF_TAPS_IN)) == (RF_INSIDE | RF_TAPS_IN))
								sample_rows<BIN, false>(p, s, cp, lx, ly, cpx, cpy, ra, r_beg, r_end, ccc, col_ok, XCx, XCy, outp, xsteps, myhist);
							else
								sample_rows<BIN, true>(p, s, cp, lx, ly, cpx, cpy, ra, r_beg, r_end, ccc, col_ok, XCx, XCy, outp, xsteps, myhist);
						}
					}
				}
			}
		}

		if (BIN == OG_BIN_OTSU) {
			// ---- the line's threshold from its histogram, then the strip is binarized in place: the CTA
			// reads back its own bytes (L2), (pixels > t) * 255   (binarize.py:11-15)
			__syncthreads();
			for (int k = tid; k < 256; k += FT) {
				unsigned int sum = 0;
#pragma unroll
				for (int q = 0; q < FHREP; ++q) {
					sum += s.hist[k * FHREP + q];
					s.hist[k * FHREP + q] = 0;
				}
				s.hsum[k] = sum;
			}
			__syncthreads();
			if (tid < 32) {
				const int t = og_otsu_from_hist_warp(s.hsum);
				if (tid == 0) {
					s.thr = t;
					if (p.thresholds != nullptr) p.thresholds[li] = t;
				}
			}
			__syncthreads();
			const unsigned int t = (unsigned int)s.thr;
			const uint8_t *g = gline;
			uint8_t *o = p.out + pix0;
			const long long head = min(npix, (long long)((16 - (reinterpret_cast<uintptr_t>(g) & 15)) & 15));
			const long long nvec = (npix - head) / 16;
			for (long long i = tid; i < head; i += FT) o[i] = g[i] > t ? 255 : 0;
			const uint4 *gv = reinterpret_cast<const uint4 *>(g + head);
			const bool out_al = ((reinterpret_cast<uintptr_t>(o + head)) & 15) == 0;
			const og_gt4_t gt = og_gt4_consts(t);
			for (long long i = tid; i < nvec; i += FT) {
				const uint4 v = gv[i];  // plain loads: these bytes were written by this CTA
				uint4 m;
				m.x = og_gt4(v.x, gt);
				m.y = og_gt4(v.y, gt);
				m.z = og_gt4(v.z, gt);
				m.w = og_gt4(v.w, gt);
				if (out_al) {
					reinterpret_cast<uint4 *>(o + head)[i] = m;
				} else {
					uint8_t *ob = o + head + i * 16;
					const unsigned int wds[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
					for (int q = 0; q < 16; ++q) ob[q] = (uint8_t)(wds[q >> 2] >> (8 * (q & 3)));
				}
			}
			for (long long i = head + nvec * 16 + tid; i < npix; i += FT) o[i] = g[i] > t ? 255 : 0;
		}
	}
}

}  // namespace

int og_launch_strips_fused(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell,
	const og_line *lines, int n, const int32_t *steps, const int64_t *pix_offsets, int32_t *chunk_offsets,
	uint8_t *out, uint8_t *gray, int64_t capacity, int otsu, int32_t *thresholds_out, cudaStream_t st) {
	FusedParams p{};
	p.src = src; p.H = H; p.W = W; p.pitch = (int)src_pitch;
	p.grid = reinterpret_cast<const float2 *>(grid);
	p.gh = gh; p.gw = gw; p.cell = cell;
	p.lines = lines; p.n = n;
	p.steps = steps; p.pix_offsets = pix_offsets;
	p.ticket = chunk_offsets + n + 1;
	p.out = out; p.gray = gray != nullptr ? gray : out;
	p.capacity = capacity;
	p.thresholds = thresholds_out;
	p.W1 = W - 1; p.H1 = H - 1;
	p.Hd = gh * cell; p.Wd = gw * cell;
	p.inv_cell = 1.0 / (double)cell;
	p.g_scale = 1.01f / (float)cell;
	cudaError_t e = cudaMemsetAsync(p.ticket, 0, sizeof(int), st);
	if (e != cudaSuccess) return og_set_error((int)e, "cudaMemsetAsync: %s", cudaGetErrorString(e));
	const int smem = (int)sizeof(FusedSmem);
	auto kern = otsu ? og_strips_fused_kernel<OG_BIN_OTSU> : og_strips_fused_kernel<OG_BIN_NONE>;
	if (og_first_use(otsu ? 1 : 0)) {  // once per device and kernel
		e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
		if (e != cudaSuccess) return og_set_error((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
	}
	const int ctas = min(n, og_num_sms() * OG_F_MINB);
	kern<<<ctas, FT, smem, st>>>(p);
	return og_check_launch("og_strips_fused_kernel");
}

extern "C" int og_extract_lines(const uint8_t *src, int H, int W, int64_t src_pitch,
	const float *grid, int gh, int gw, int cell, const float *map,
	const og_line *lines, int n,
	const int32_t *steps, const int64_t *pix_offsets, int32_t *chunk_offsets,
	uint8_t *strips_out, int64_t capacity, int32_t *bbox_out,
	int binarize, int window, uint8_t *gray_out, int32_t *thresholds_out, void *stream) {
	const bool gather = (binarize & OG_FLAG_MAP_GATHER) != 0;
	binarize &= ~OG_FLAG_MAP_GATHER;
	OG_REQUIRE(binarize == OG_BIN_NONE || binarize == OG_BIN_OTSU || binarize == OG_BIN_SAUVOLA, OG_E_ARG,
		"binarize %d unsupported", binarize);
	if (binarize == OG_BIN_SAUVOLA) {
		OG_REQUIRE(window >= 3 && (window & 1) == 1, OG_E_WINDOW, "Window size w = %d must be odd and greater than 1.", window);
		OG_REQUIRE(gray_out != nullptr && gray_out != strips_out, OG_E_ARG, "Sauvola needs a separate gray_out buffer of `capacity` bytes");
	}
	cudaStream_t st = (cudaStream_t)stream;
	// Engine choice (measured on B200, 2000 lines of a 4800x7200 page, profiles/README.md): with the
	// materialised map the screened map-gather engine of og_lines.cu is the fastest (0.217 ms, Otsu as
	// a second pass over the L2-resident strips 0.042 ms); without a map the line-owning engine below
	// recomputes the nodes on chip (0.37 ms; Otsu fused, +0.06 ms) and reads no dense map at all.
	if (bbox_out != nullptr || gather || map != nullptr) {
		uint8_t *target = strips_out;
		if (binarize == OG_BIN_OTSU && gray_out != nullptr) target = gray_out;  // else Otsu runs in place
		if (binarize == OG_BIN_SAUVOLA) target = gray_out;
		int rc = og_extract_lines_engines(src, H, W, src_pitch, grid, gh, gw, cell, map, lines, n, steps, pix_offsets,
			chunk_offsets, target, capacity, bbox_out, st);
		if (rc || n == 0 || binarize == OG_BIN_NONE) return rc;
		return og_binarize_strips_impl(target, steps, pix_offsets, chunk_offsets, n, binarize, window, strips_out, capacity,
			thresholds_out, st);
	}
	OG_REQUIRE(grid != nullptr, OG_E_ARG, "grid is NULL");
	OG_REQUIRE(gh >= 2 && gw >= 2 && cell >= 1, OG_E_ARG, "grid shape (%d, %d) cell %d invalid", gh, gw, cell);
	OG_REQUIRE(((reinterpret_cast<uintptr_t>(grid) | reinterpret_cast<uintptr_t>(lines)) & 7) == 0,
		OG_E_ALIGN, "grid and lines must be 8-byte aligned");
	OG_REQUIRE((int64_t)gh * cell < 32767 && (int64_t)gw * cell < 32767, OG_E_RANGE, "dewarped size exceeds OpenCV's 32767 limit");
	OG_REQUIRE(n >= 0, OG_E_ARG, "n < 0");
	if (n == 0) return OG_OK;
	OG_REQUIRE(lines != nullptr && steps != nullptr && chunk_offsets != nullptr && pix_offsets != nullptr && strips_out != nullptr,
		OG_E_ARG, "lines / steps / offsets / strips_out is NULL");
	OG_REQUIRE(src != nullptr && H > 0 && W > 0 && src_pitch >= W, OG_E_ARG, "src invalid");
	OG_REQUIRE(H < 32767 && W < 32767, OG_E_RANGE, "source %d x %d exceeds OpenCV's 32767 limit", W, H);
	OG_REQUIRE((int64_t)H * src_pitch < ((int64_t)1 << 31), OG_E_RANGE, "source image larger than 2 GiB");
	if (binarize == OG_BIN_SAUVOLA) {
		int rc = og_launch_strips_fused(src, H, W, src_pitch, grid, gh, gw, cell, lines, n, steps, pix_offsets, chunk_offsets,
			gray_out, nullptr, capacity, 0, nullptr, st);
		if (rc) return rc;
		return og_binarize_strips_impl(gray_out, steps, pix_offsets, chunk_offsets, n, binarize, window, strips_out, capacity, nullptr, st);
	}
	return og_launch_strips_fused(src, H, W, src_pitch, grid, gh, gw, cell, lines, n, steps, pix_offsets, chunk_offsets,
		strips_out, gray_out, capacity, binarize == OG_BIN_OTSU, thresholds_out, st);
}
