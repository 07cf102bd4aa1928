/*
 * oracle/sad_oracle.c -- CPU restatement of the displacement-search hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call it, and there only as the checker.
 *
 * PARITY STATUS: "parity unpinned".  The reference (ruzko/motion-estimation) has
 * no block-matching search at all: its per-pair displacement is
 * cv.findTransformECC (estimate_vel_vid.py:306-369, estimate_vel.py:286-297) and
 * it ships no tests or golden vectors.  This file therefore DEFINES the integer
 * search the CUDA path must reproduce bit-exactly; what ties it back to the
 * reference is (a) the call-site contract -- two uint8 (H,W) luma frames in, one
 * (pdx, pdy, flag) out, +x content motion => pdx > 0 (estimate_vel_vid.py:522,
 * :358) -- and (b) the soft checks in tests/ against oracle/ecc_port.py (the
 * restated ECC driver run on cv2) and the nominal mm/min of the example clips.
 *
 * The frozen specification ("the spec"):
 *
 *   grid      non-overlapping BxB blocks, nbx = W / B, nby = H / B (floor; the
 *             ragged right/bottom strip is not searched).  Block (bx,by) has its
 *             top-left pixel at (x0,y0) = (bx*B, by*B) in the PREVIOUS frame.
 *   cost      SAD(dx,dy) = sum_{i,j<B} | cur[y0+dy+i][x0+dx+j] - prev[y0+i][x0+j] |
 *   domain    dx in [-R,R], dy in [-Ry,Ry] (Ry = R unless range_y names a smaller vertical
 *             range: the recordings move along x only) and the displaced block must lie wholly inside the
 *             frame: 0 <= x0+dx, x0+dx+B <= W, 0 <= y0+dy, y0+dy+B <= H.
 *             (0,0) is always in the domain.  No padding pixels are invented.
 *   argmin    first minimum in raster order, dy outer ascending, dx inner
 *             ascending, strict '<' -- i.e. numpy.argmin over a C-ordered
 *             (2R+1, 2R+1) array indexed [dy+R][dx+R] with out-of-domain = +inf.
 *   textured  flag bit 0: max - min of the block's PREVIOUS-frame pixels
 *             >= min_range.
 *   interior  per axis.  flag bit 1 (X): (dx-1,dy) and (dx+1,dy) are both in the
 *             domain; flag bit 2 (Y): (dx,dy-1) and (dx,dy+1) are.  (Per axis
 *             because the example recordings are 64 px tall: the top and bottom
 *             block rows can never be interior in y, yet measure x perfectly.)
 *   peaked    per axis, with c0 the minimum and cm / cp the costs one step below
 *             / above on that axis:  den = cm - 2*c0 + cp  (exact integer, >= 0).
 *             flag bit 3 (X) / bit 4 (Y): interior on that axis && den >= min_curv
 *             (min_curv >= 1, so a peaked axis has a strict parabola).
 *   subpixel  per axis, only when that axis is interior, den > 0 and
 *             p->subpixel != 0:
 *               off = 0.5f * ((float)(cm - cp) / (float)den)
 *             in IEEE binary32, one correctly rounded divide, then one exact
 *             multiply.  sub = (float)d + off.  |off| <= 0.5 by construction.
 *             Otherwise sub = (float)d.
 *   valid     per axis: textured && peaked(axis).
 *   reduce    per axis, the median over that axis' valid blocks of sub, defined
 *             as 0.5f * (s[(n-1)/2] + s[n/2]) on the ascending sort s (binary32,
 *             add then multiply).  nvalid_x / nvalid_y = the two counts.
 *             flag = nvalid_x >= max(min_valid,1)  (x is the feed axis of the
 *             recordings; the reference gates on pdx, estimate_vel_vid.py:531).
 *             pdx = 0 when nvalid_x == 0, pdy = 0 when nvalid_y == 0.
 *   mean      per axis, over the same valid blocks: S = sum of llrintf(sub * 2^20)
 *             (exact 64-bit integer, order independent), mean = (float)((double)S /
 *             ((double)n * 1048576.0)); 0 when n == 0.  This is the frame mean of the
 *             dense per-block field, the role the reference gives its Farneback flow
 *             (estimate_vel_vid.py:385-391).
 *             n_textured = number of blocks with the textured bit: nvalid_x /
 *             n_textured is the share of usable blocks whose minimum is a strict
 *             interior peak, a confidence figure for the pair (it drops when the
 *             motion leaves the search range).
 *
 * Build: see oracle/Makefile (gcc -O3 -fopenmp -shared -fPIC).
 */
#include <immintrin.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <limits.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ME_FLAG_TEXTURED   1u
#define ME_FLAG_X_INTERIOR 2u
#define ME_FLAG_Y_INTERIOR 4u
#define ME_FLAG_X_PEAKED   8u
#define ME_FLAG_Y_PEAKED   16u

typedef struct {
    int32_t block;      /* B: 4..64, multiple of 4 */
    int32_t range;      /* R: 1..128 */
    int32_t min_range;  /* texture gate, in grey levels */
    int32_t subpixel;   /* 0 = integer only */
    int32_t min_valid;  /* pair accepted when nvalid_x >= max(min_valid,1) */
    int32_t min_curv;   /* peak gate on den = cm - 2 c0 + cp; values < 1 are treated as 1 */
    int32_t range_y;    /* Ry: vertical range, 1..R; 0 = R (square search) */
} me_oracle_params;

/* Plain statement of the cost (what the spec says), used for block sizes the SIMD form below does not cover and,
 * with ME_ORACLE_SCALAR defined, everywhere: tests/test_oracle.py cross-checks the NumPy restatement either way. */
static inline uint32_t block_sad_scalar(const uint8_t *a, size_t pa, const uint8_t *b, size_t pb, int B)
{
    uint32_t s = 0;
    for (int i = 0; i < B; ++i) {
        const uint8_t *ra = a + (size_t)i * pa, *rb = b + (size_t)i * pb;
        uint32_t r = 0;
        for (int j = 0; j < B; ++j) r += (uint32_t)abs((int)ra[j] - (int)rb[j]);
        s += r;
    }
    return s;
}

/* The same integer, computed with psadbw (8 or 16 absolute differences summed per instruction) where the block
 * width allows: this is what a CPU implementation of the search would use, so the timed CPU baseline of bench.py
 * is not a straw man (round 1's auto-vectorised scalar loop did 139 G cand*px/s on 16 cores). */
static inline uint32_t block_sad(const uint8_t *a, size_t pa, const uint8_t *b, size_t pb, int B)
{
#if defined(__SSE2__) && !defined(ME_ORACLE_SCALAR)
    if ((B & 15) == 0) {
        __m128i acc = _mm_setzero_si128();
        for (int i = 0; i < B; ++i)
            for (int j = 0; j < B; j += 16)
                acc = _mm_add_epi64(acc, _mm_sad_epu8(_mm_loadu_si128((const __m128i *)(a + (size_t)i * pa + j)),
                                                      _mm_loadu_si128((const __m128i *)(b + (size_t)i * pb + j))));
        return (uint32_t)(_mm_cvtsi128_si64(acc) + _mm_cvtsi128_si64(_mm_unpackhi_epi64(acc, acc)));
    }
    if (B == 8) {
        __m128i acc = _mm_setzero_si128();
        for (int i = 0; i < 8; ++i)
            acc = _mm_add_epi64(acc, _mm_sad_epu8(_mm_loadl_epi64((const __m128i *)(a + (size_t)i * pa)),
                                                  _mm_loadl_epi64((const __m128i *)(b + (size_t)i * pb))));
        return (uint32_t)_mm_cvtsi128_si64(acc);
    }
#endif
    return block_sad_scalar(a, pa, b, pb, B);
}

static int cmp_float(const void *a, const void *b)
{
    float x = *(const float *)a, y = *(const float *)b;
    return (x > y) - (x < y);
}

/* median as the spec defines it; v is clobbered (sorted). n >= 1. */
/* spec: mean -- a refined displacement as a 2^-20 fixed-point integer (the product is exact in binary32, the
 * rounding is to nearest even), so that the sum over blocks is exact and independent of the order of addition */
static int64_t spec_fixed(float v) { return (int64_t)llrintf(v * 1048576.0f); }

static float spec_median(float *v, int n)
{
    qsort(v, (size_t)n, sizeof(float), cmp_float);
    float lo = v[(n - 1) / 2], hi = v[n / 2];
    float s = lo + hi;
    return 0.5f * s;
}

int me_oracle_num_blocks(int H, int W, int B, int *nbx, int *nby)
{
    if (B <= 0) return -1;
    *nbx = W / B;
    *nby = H / B;
    return 0;
}

/* Cost surface of one block: out[(dy+R)*(2R+1) + (dx+R)], UINT32_MAX outside the domain. */
int me_oracle_cost_surface(const uint8_t *prev, const uint8_t *cur, int H, int W, size_t pitch,
                           int B, int R, int bx, int by, uint32_t *out)
{
    int x0 = bx * B, y0 = by * B, U = 2 * R + 1;
    if (x0 + B > W || y0 + B > H || bx < 0 || by < 0) return -1;
    for (int dy = -R; dy <= R; ++dy)
        for (int dx = -R; dx <= R; ++dx) {
            uint32_t c = UINT32_MAX;
            if (x0 + dx >= 0 && x0 + dx + B <= W && y0 + dy >= 0 && y0 + dy + B <= H)
                c = block_sad(prev + (size_t)y0 * pitch + x0, pitch,
                              cur + (size_t)(y0 + dy) * pitch + (x0 + dx), pitch, B);
            out[(size_t)(dy + R) * U + (dx + R)] = c;
        }
    return 0;
}

/*
 * One frame pair.  Per-block outputs are row-major over (by,bx):
 *   blk_mv    int16  [nb][2]  (dx,dy)
 *   blk_sad   uint32 [nb]     minimum SAD
 *   blk_sub   float  [nb][2]  refined (dx,dy)
 *   blk_flags uint8  [nb]
 * Per-pair outputs: pair_d[4] = (pdx, pdy, mean_dx, mean_dy) in full-resolution pixels,
 *                   pair_i[4] = (flag, nvalid_x, nvalid_y, n_textured).
 * Any per-block pointer may be NULL.  Returns 0, or <0 on bad arguments.
 */
int me_oracle_search(const uint8_t *prev, const uint8_t *cur, int H, int W, size_t pitch,
                     const me_oracle_params *p, int16_t *blk_mv, uint32_t *blk_sad, float *blk_sub,
                     uint8_t *blk_flags, float *pair_d, int32_t *pair_i, int nthreads)
{
    const int B = p->block, R = p->range, Ry = p->range_y > 0 ? p->range_y : p->range;
    if (B < 4 || B > 64 || (B & 3) || R < 1 || R > 128 || Ry > R || H < B || W < B || pitch < (size_t)W) return -1;
    const int nbx = W / B, nby = H / B, nb = nbx * nby;
    const int32_t min_curv = p->min_curv < 1 ? 1 : p->min_curv;
    float *sx = (float *)malloc(sizeof(float) * (size_t)nb * 2);
    uint8_t *fl = (uint8_t *)malloc((size_t)nb);
    if (!sx || !fl) { free(sx); free(fl); return -2; }
    float *sy = sx + nb;
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
#pragma omp parallel for schedule(dynamic, 4)
    for (int b = 0; b < nb; ++b) {
        const int bx = b % nbx, by = b / nbx, x0 = bx * B, y0 = by * B;
        const uint8_t *ref = prev + (size_t)y0 * pitch + x0;
        /* domain (an axis-aligned box of displacements) */
        const int dxlo = (-R > -x0) ? -R : -x0, dxhi = (R < W - B - x0) ? R : W - B - x0;
        const int dylo = (-Ry > -y0) ? -Ry : -y0, dyhi = (Ry < H - B - y0) ? Ry : H - B - y0;
        uint32_t best = UINT32_MAX;
        int bdx = 0, bdy = 0;
        for (int dy = dylo; dy <= dyhi; ++dy)
            for (int dx = dxlo; dx <= dxhi; ++dx) {
                uint32_t c = block_sad(ref, pitch, cur + (size_t)(y0 + dy) * pitch + (x0 + dx), pitch, B);
                if (c < best) { best = c; bdx = dx; bdy = dy; }
            }
        unsigned flags = 0;
        int mn = 255, mx = 0;
        for (int i = 0; i < B; ++i)
            for (int j = 0; j < B; ++j) {
                int v = ref[(size_t)i * pitch + j];
                if (v < mn) mn = v;
                if (v > mx) mx = v;
            }
        if (mx - mn >= p->min_range) flags |= ME_FLAG_TEXTURED;
        float fx = (float)bdx, fy = (float)bdy;
        const uint8_t *c0p = cur + (size_t)(y0 + bdy) * pitch + (x0 + bdx);
        const int32_t c0 = (int32_t)best;
        if (bdx - 1 >= dxlo && bdx + 1 <= dxhi) {
            flags |= ME_FLAG_X_INTERIOR;
            int32_t cm = (int32_t)block_sad(ref, pitch, c0p - 1, pitch, B);
            int32_t cp = (int32_t)block_sad(ref, pitch, c0p + 1, pitch, B);
            int32_t den = cm - 2 * c0 + cp;
            if (den >= min_curv) flags |= ME_FLAG_X_PEAKED;
            if (den > 0 && p->subpixel) { float q = (float)(cm - cp) / (float)den; fx = fx + 0.5f * q; }
        }
        if (bdy - 1 >= dylo && bdy + 1 <= dyhi) {
            flags |= ME_FLAG_Y_INTERIOR;
            int32_t cm = (int32_t)block_sad(ref, pitch, c0p - pitch, pitch, B);
            int32_t cp = (int32_t)block_sad(ref, pitch, c0p + pitch, pitch, B);
            int32_t den = cm - 2 * c0 + cp;
            if (den >= min_curv) flags |= ME_FLAG_Y_PEAKED;
            if (den > 0 && p->subpixel) { float q = (float)(cm - cp) / (float)den; fy = fy + 0.5f * q; }
        }
        if (blk_mv) { blk_mv[2 * b] = (int16_t)bdx; blk_mv[2 * b + 1] = (int16_t)bdy; }
        if (blk_sad) blk_sad[b] = best;
        if (blk_sub) { blk_sub[2 * b] = fx; blk_sub[2 * b + 1] = fy; }
        if (blk_flags) blk_flags[b] = (uint8_t)flags;
        sx[b] = fx; sy[b] = fy;
        fl[b] = (uint8_t)flags;
    }
    int nx = 0, ny = 0, ntex = 0;
    int64_t fsx = 0, fsy = 0;        /* fixed-point sums of the valid refined displacements (spec: mean) */
    for (int b = 0; b < nb; ++b) {   /* in-place compaction: write index never passes read index */
        float vx = sx[b], vy = sy[b];
        if (fl[b] & ME_FLAG_TEXTURED) ++ntex;
        if ((fl[b] & (ME_FLAG_TEXTURED | ME_FLAG_X_PEAKED)) == (ME_FLAG_TEXTURED | ME_FLAG_X_PEAKED)) { sx[nx++] = vx; fsx += spec_fixed(vx); }
        if ((fl[b] & (ME_FLAG_TEXTURED | ME_FLAG_Y_PEAKED)) == (ME_FLAG_TEXTURED | ME_FLAG_Y_PEAKED)) { sy[ny++] = vy; fsy += spec_fixed(vy); }
    }
    int need = p->min_valid > 1 ? p->min_valid : 1;
    pair_d[0] = nx > 0 ? spec_median(sx, nx) : 0.0f;
    pair_d[1] = ny > 0 ? spec_median(sy, ny) : 0.0f;
    pair_d[2] = nx > 0 ? (float)((double)fsx / ((double)nx * 1048576.0)) : 0.0f;
    pair_d[3] = ny > 0 ? (float)((double)fsy / ((double)ny * 1048576.0)) : 0.0f;
    pair_i[0] = nx >= need;
    pair_i[1] = nx;
    pair_i[2] = ny;
    pair_i[3] = ntex;
    free(sx); free(fl);
    return 0;
}

/*
 * Batch form used for timing and for the anchored pair loop: frames is (N,H,pitch),
 * pairs is P x (prev_index, cur_index).  Per-pair outputs only: pair_d (P,4), pair_i (P,4).
 */
int me_oracle_search_pairs(const uint8_t *frames, int N, int H, int W, size_t pitch,
                           const me_oracle_params *p, const int32_t *pairs, int P,
                           float *pair_d, int32_t *pair_i, int nthreads)
{
    for (int k = 0; k < P; ++k) {
        int a = pairs[2 * k], c = pairs[2 * k + 1];
        if (a < 0 || a >= N || c < 0 || c >= N) return -3;
        int rc = me_oracle_search(frames + (size_t)a * H * pitch, frames + (size_t)c * H * pitch, H, W,
                                  pitch, p, NULL, NULL, NULL, NULL, pair_d + 4 * k, pair_i + 4 * k, nthreads);
        if (rc) return rc;
    }
    return 0;
}

/*
 * BGR -> luma exactly as cv.cvtColor(COLOR_BGR2GRAY) does it for 8-bit input
 * (reference call sites estimate_vel_vid.py:148,178): 15-bit fixed point,
 * Y = (B*3735 + G*19235 + R*9798 + 16384) >> 15.  Pinned against the installed cv2 (4.13) in
 * tests/test_oracle.py; the reference's own pin is opencv 4.6.0.66, whose coefficients cannot be
 * checked offline.
 */
int me_oracle_bgr2gray(const uint8_t *bgr, size_t npix, uint8_t *gray)
{
    for (size_t i = 0; i < npix; ++i) {
        uint32_t b = bgr[3 * i], g = bgr[3 * i + 1], r = bgr[3 * i + 2];
        gray[i] = (uint8_t)((b * 3735u + g * 19235u + r * 9798u + 16384u) >> 15);
    }
    return 0;
}

int me_oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
