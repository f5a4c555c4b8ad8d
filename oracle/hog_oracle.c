/* oracle/hog_oracle.c -- TEST INFRASTRUCTURE ONLY (see hog_oracle.h).
 *
 * Plain-C restatement of the reference's CPU HOG path.  Compile with -ffp-contract=off so that
 * a*b+c is never fused: the reference is built for SSE3 (no FMA) and the float results here are meant
 * to be EQUAL to its results, not merely close.
 */
#include "hog_oracle.h"
#include <math.h>
#include <float.h>
#include <stdlib.h>
#include <string.h>

/* cvRound: round-half-to-even (core/include/opencv2/core/fast_math.hpp, SSE2 cvtsd2si path). */
static int cv_round(double v) { return (int)lrint(v); }
static int cv_roundf(float v) { return (int)lrintf(v); }
static int cv_floorf(float v) { int i = (int)v; return i - (v < (float)i); }
static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

void hogo_default_geom(hogo_geom* g)
{
    g->win_w = 64; g->win_h = 128; g->block_w = 16; g->block_h = 16;
    g->stride_w = 8; g->stride_h = 8; g->cell_w = 8; g->cell_h = 8; g->nbins = 9;
    g->win_sigma = -1.0; g->l2hys = 0.2; g->gamma = 1;
}

/* ------------------------------------------------------------------ levels */
/* cudaobjdetect/src/hog.cpp:424-441: scale accumulated as a product, cvRound in the fit test, the level
 * that fails the test is dropped, at least one level kept. */
int hogo_levels(int W, int H, int win_w, int win_h, double scale0, int nlevels,
                double* scales, int* widths, int* heights, int cap)
{
    double scale = 1.0;
    int levels = 0, n = 0;
    for (levels = 0; levels < nlevels; levels++) {
        if (n < cap) scales[n] = scale;
        n++;
        if (cv_round(W / scale) < win_w || cv_round(H / scale) < win_h || scale0 <= 1)
            break;
        scale *= scale0;
    }
    levels = imax(levels, 1);
    if (levels > cap) levels = cap;
    for (int i = 0; i < levels; i++) {
        widths[i] = cv_round(W / scales[i]);
        heights[i] = cv_round(H / scales[i]);
    }
    return levels;
}

/* ------------------------------------------------------------------ resize */
static int clipi(int x, int a, int b) { return x >= a ? (x < b ? x : b - 1) : a; }

void hogo_resize_linear(const unsigned char* src, int sw, int sh, int cn, long spitch,
                        unsigned char* dst, int dw, int dh, long dpitch)
{
    /* resize.cpp:3745-3750 */
    if (dw == sw && dh == sh) {
        for (int y = 0; y < sh; y++) memcpy(dst + y * dpitch, src + y * spitch, (size_t)sw * cn);
        return;
    }
    /* resize.cpp:3730-3731, 3481-3489 */
    double inv_scale_x = (double)dw / sw, inv_scale_y = (double)dh / sh;
    double scale_x = 1. / inv_scale_x, scale_y = 1. / inv_scale_y;
    int iscale_x = (int)lrint(scale_x), iscale_y = (int)lrint(scale_y);
    int is_area_fast = fabs(scale_x - iscale_x) < DBL_EPSILON && fabs(scale_y - iscale_y) < DBL_EPSILON;

    /* resize.cpp:3523-3554 + 2542-2578: exact 2x decimation is INTER_AREA */
    if (is_area_fast && iscale_x == 2 && iscale_y == 2) {
        for (int y = 0; y < dh; y++) {
            const unsigned char* S = src + (long)(2 * y) * spitch;
            const unsigned char* N = S + spitch;
            unsigned char* D = dst + y * dpitch;
            for (int x = 0; x < dw; x++)
                for (int k = 0; k < cn; k++) {
                    int i0 = 2 * x * cn + k;
                    D[x * cn + k] = (unsigned char)((S[i0] + S[i0 + cn] + N[i0] + N[i0 + cn] + 2) >> 2);
                }
        }
        return;
    }

    /* resize.cpp:3609-3696: coefficient tables */
    int* xofs = (int*)malloc(sizeof(int) * (size_t)dw);
    short* ialpha = (short*)malloc(sizeof(short) * 2 * (size_t)dw);
    int* yofs = (int*)malloc(sizeof(int) * (size_t)dh);
    short* ibeta = (short*)malloc(sizeof(short) * 2 * (size_t)dh);
    int xmax = dw;
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx + 1 >= sw) {
            xmax = imin(xmax, dx);
            if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        }
        xofs[dx] = sx;
        float c0 = 1.f - fx, c1 = fx;
        ialpha[2 * dx] = (short)cv_roundf(c0 * 2048);
        ialpha[2 * dx + 1] = (short)cv_roundf(c1 * 2048);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floorf(fy);
        fy -= sy;
        yofs[dy] = sy;
        float c0 = 1.f - fy, c1 = fy;
        ibeta[2 * dy] = (short)cv_roundf(c0 * 2048);
        ibeta[2 * dy + 1] = (short)cv_roundf(c1 * 2048);
    }

    /* resize.cpp:1491-1542 (horizontal, int accumulators), 1780-1857 (row clipping),
     * 1577-1603 (vertical, two truncating shifts) */
    int* row0 = (int*)malloc(sizeof(int) * (size_t)dw * cn);
    int* row1 = (int*)malloc(sizeof(int) * (size_t)dw * cn);
    for (int dy = 0; dy < dh; dy++) {
        int sy0 = clipi(yofs[dy], 0, sh), sy1 = clipi(yofs[dy] + 1, 0, sh);
        const unsigned char* S0 = src + (long)sy0 * spitch;
        const unsigned char* S1 = src + (long)sy1 * spitch;
        for (int dx = 0; dx < dw; dx++) {
            int sx = xofs[dx] * cn;
            int a0 = ialpha[2 * dx], a1 = ialpha[2 * dx + 1];
            for (int k = 0; k < cn; k++) {
                if (dx < xmax) {
                    row0[dx * cn + k] = S0[sx + k] * a0 + S0[sx + cn + k] * a1;
                    row1[dx * cn + k] = S1[sx + k] * a0 + S1[sx + cn + k] * a1;
                } else {
                    row0[dx * cn + k] = S0[sx + k] * 2048;
                    row1[dx * cn + k] = S1[sx + k] * 2048;
                }
            }
        }
        int b0 = ibeta[2 * dy], b1 = ibeta[2 * dy + 1];
        unsigned char* D = dst + dy * dpitch;
        for (int x = 0; x < dw * cn; x++)
            D[x] = (unsigned char)((((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2);
    }
    free(row0); free(row1); free(xofs); free(ialpha); free(yofs); free(ibeta);
}

/* ------------------------------------------------------------------ gradient */
/* core/src/mathfuncs_core.simd.hpp:45-71 (scalar) == :92-103 (SIMD): degree-7 odd polynomial, degrees. */
static float fast_atan_deg(float y, float x)
{
    static const float p1 = 0.9997878412794807f * (float)(180 / 3.1415926535897932384626433832795);
    static const float p3 = -0.3258083974640975f * (float)(180 / 3.1415926535897932384626433832795);
    static const float p5 = 0.1555786518463281f * (float)(180 / 3.1415926535897932384626433832795);
    static const float p7 = -0.04432655554792128f * (float)(180 / 3.1415926535897932384626433832795);
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

static int reflect101(int p, int len)
{
    if (len == 1) return 0;
    while (p < 0 || p >= len) { if (p < 0) p = -p; else p = 2 * len - 2 - p; }
    return p;
}

void hogo_compute_gradient(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                           float* grad, unsigned char* qangle)
{
    float lut[256];
    for (int i = 0; i < 256; i++) lut[i] = g->gamma ? sqrtf((float)i) : (float)i;  /* hog.cpp:258-300 */
    const int nbins = g->nbins;
    const float angleScale = (float)(nbins / 3.1415926535897932384626433832795);   /* hog.cpp:345 */
    const float rad = (float)(3.1415926535897932384626433832795 / 180);            /* mathfuncs_core:125 */
    const int ncol = cn == 1 ? 1 : 3;  /* channel 3 (alpha) never read */
    float* dxb = (float*)malloc(sizeof(float) * 2 * (size_t)w);
    float* dyb = dxb + w;
    for (int y = 0; y < h; y++) {
        const unsigned char* cur = img + (long)y * pitch;
        const unsigned char* prev = img + (long)reflect101(y - 1, h) * pitch;
        const unsigned char* next = img + (long)reflect101(y + 1, h) * pitch;
        for (int x = 0; x < w; x++) {
            int xl = reflect101(x - 1, w) * cn, xr = reflect101(x + 1, w) * cn, xc = x * cn;
            if (ncol == 1) {                                       /* hog.cpp:357-365 */
                dxb[x] = lut[cur[xr]] - lut[cur[xl]];
                dyb[x] = lut[next[xc]] - lut[prev[xc]];
            } else {
                float dx0 = lut[cur[xr]] - lut[cur[xl]],         dy0 = lut[next[xc]] - lut[prev[xc]];
                float dx1 = lut[cur[xr + 1]] - lut[cur[xl + 1]], dy1 = lut[next[xc + 1]] - lut[prev[xc + 1]];
                float dx2 = lut[cur[xr + 2]] - lut[cur[xl + 2]], dy2 = lut[next[xc + 2]] - lut[prev[xc + 2]];
                float m0 = dx0 * dx0 + dy0 * dy0, m1 = dx1 * dx1 + dy1 * dy1, m2 = dx2 * dx2 + dy2 * dy2;
                float dx, dy;
                if (x < (w & ~3)) {
                    /* SSE2 body, hog.cpp:369-407: ch2 only if strictly larger than ch1; the winner only if
                     * strictly larger than ch0 -> lowest channel wins ties */
                    if (m2 > m1) { dx = dx2; dy = dy2; } else { dx = dx1; dy = dy1; }
                    float mm = m2 > m1 ? m2 : m1;
                    if (!(mm > m0)) { dx = dx0; dy = dy0; }
                } else {
                    /* scalar tail (last w%4 columns), hog.cpp:449-479: starts from ch2, replaced only by a
                     * strictly larger one -> highest channel wins ties */
                    float mag0 = m2; dx = dx2; dy = dy2;
                    if (mag0 < m1) { dx = dx1; dy = dy1; mag0 = m1; }
                    if (mag0 < m0) { dx = dx0; dy = dy0; mag0 = m0; }
                }
                dxb[x] = dx; dyb[x] = dy;
            }
        }
        for (int x = 0; x < w; x++) {
            float dx = dxb[x], dy = dyb[x];
            float mag = sqrtf(dx * dx + dy * dy);                 /* mathfuncs_core:194-225 */
            float ang = fast_atan_deg(dy, dx) * rad;              /* mathfuncs_core:123-159 */
            float angle = ang * angleScale - 0.5f;                /* hog.cpp:486-547 */
            int hidx = cv_floorf(angle);
            angle -= hidx;
            grad[((size_t)y * w + x) * 2] = mag * (1.f - angle);
            grad[((size_t)y * w + x) * 2 + 1] = mag * angle;
            if (hidx < 0) hidx += nbins; else if (hidx >= nbins) hidx -= nbins;
            qangle[((size_t)y * w + x) * 2] = (unsigned char)hidx;
            hidx++;
            if (hidx >= nbins) hidx = 0;
            qangle[((size_t)y * w + x) * 2 + 1] = (unsigned char)hidx;
        }
    }
    free(dxb);
}

/* ------------------------------------------------------------------ block histograms */
typedef struct { int i, j, n; int ofs[4]; float hw[4]; float gw; } pixdata;

/* HOGCache::init, hog.cpp:653-839: Gaussian block window and per-pixel cell interpolation table, kept in
 * the reference's processing order: 1-cell pixels, then 2-cell, then 4-cell, each scanned j-outer/i-inner. */
static int build_pixdata(const hogo_geom* g, pixdata* out)
{
    const int bw = g->block_w, bh = g->block_h, nbins = g->nbins;
    const int ncx = bw / g->cell_w, ncy = bh / g->cell_h;
    float sigma = (float)(g->win_sigma > 0 ? g->win_sigma : (bw + bh) / 8.);   /* hog.cpp:101-104 */
    float scale = 1.f / (sigma * sigma * 2);
    int n = 0;
    for (int pass = 1; pass <= 4; pass *= 2)
        for (int j = 0; j < bw; j++)
            for (int i = 0; i < bh; i++) {
                pixdata d; memset(&d, 0, sizeof d);
                float cellX = (j + 0.5f) / g->cell_w - 0.5f, cellY = (i + 0.5f) / g->cell_h - 0.5f;
                int x0 = cv_floorf(cellX), y0 = cv_floorf(cellY), x1 = x0 + 1, y1 = y0 + 1;
                cellX -= x0; cellY -= y0;
                int xin0 = (unsigned)x0 < (unsigned)ncx, xin1 = (unsigned)x1 < (unsigned)ncx;
                int yin0 = (unsigned)y0 < (unsigned)ncy, yin1 = (unsigned)y1 < (unsigned)ncy;
                if (xin0 && xin1) {
                    if (yin0 && yin1) {
                        d.n = 4;
                        d.ofs[0] = (x0 * ncy + y0) * nbins; d.hw[0] = (1.f - cellX) * (1.f - cellY);
                        d.ofs[1] = (x1 * ncy + y0) * nbins; d.hw[1] = cellX * (1.f - cellY);
                        d.ofs[2] = (x0 * ncy + y1) * nbins; d.hw[2] = (1.f - cellX) * cellY;
                        d.ofs[3] = (x1 * ncy + y1) * nbins; d.hw[3] = cellX * cellY;
                    } else {
                        d.n = 2;
                        if (yin0) { y1 = y0; cellY = 1.f - cellY; }
                        d.ofs[0] = (x0 * ncy + y1) * nbins; d.hw[0] = (1.f - cellX) * cellY;
                        d.ofs[1] = (x1 * ncy + y1) * nbins; d.hw[1] = cellX * cellY;
                    }
                } else {
                    if (xin0) { x1 = x0; cellX = 1.f - cellX; }
                    if (yin0 && yin1) {
                        d.n = 2;
                        d.ofs[0] = (x1 * ncy + y0) * nbins; d.hw[0] = cellX * (1.f - cellY);
                        d.ofs[1] = (x1 * ncy + y1) * nbins; d.hw[1] = cellX * cellY;
                    } else {
                        d.n = 1;
                        if (yin0) { y1 = y0; cellY = 1.f - cellY; }
                        d.ofs[0] = (x1 * ncy + y1) * nbins; d.hw[0] = cellX * cellY;
                    }
                }
                if (d.n != pass) continue;
                d.i = i; d.j = j;
                float di = i - bh * 0.5f, dj = j - bw * 0.5f;
                di *= di; dj *= dj;
                d.gw = expf(-(di + dj) * scale);
                out[n++] = d;
            }
    return n;
}

/* HOGCache::normalizeBlockHistogram, hog.cpp:1119-1248 (SSE2 lane order). */
static void normalize_block(float* hist, int sz, float thresh)
{
    float part[4] = {0, 0, 0, 0}, sum;
    int i;
    for (i = 0; i <= sz - 4; i += 4)
        for (int l = 0; l < 4; l++) part[l] = i == 0 ? hist[l] * hist[l] : part[l] + hist[i + l] * hist[i + l];
    { float t0 = part[0] + part[1], t1 = part[2] + part[3]; sum = t0 + t1; }
    for (; i < sz; i++) sum += hist[i] * hist[i];
    float scale = 1.f / (sqrtf(sum) + sz * 0.1f);
    part[0] = part[1] = part[2] = part[3] = 0;
    for (i = 0; i <= sz - 4; i += 4)
        for (int l = 0; l < 4; l++) {
            float p = hist[i + l] * scale;
            p = p < thresh ? p : thresh;           /* _mm_min_ps(p, thr) */
            hist[i + l] = p;
            part[l] = i == 0 ? p * p : part[l] + p * p;
        }
    { float t0 = part[0] + part[1], t1 = part[2] + part[3]; sum = t0 + t1; }
    for (; i < sz; i++) {
        float p = hist[i] * scale;
        hist[i] = p < thresh ? p : thresh;
        sum += hist[i] * hist[i];
    }
    scale = 1.f / (sqrtf(sum) + 1e-3f);
    for (i = 0; i < sz; i++) hist[i] *= scale;
}

long hogo_block_hists(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                      float* out, int* blocks_x, int* blocks_y)
{
    const int bw = g->block_w, bh = g->block_h;
    int nbx = w >= bw ? (w - bw) / g->stride_w + 1 : 0;
    int nby = h >= bh ? (h - bh) / g->stride_h + 1 : 0;
    if (blocks_x) *blocks_x = nbx;
    if (blocks_y) *blocks_y = nby;
    if (nbx <= 0 || nby <= 0) return 0;
    const int hsz = (bw / g->cell_w) * (bh / g->cell_h) * g->nbins;
    float* grad = (float*)malloc(sizeof(float) * 2 * (size_t)w * h);
    unsigned char* qa = (unsigned char*)malloc(2 * (size_t)w * h);
    hogo_compute_gradient(g, img, w, h, cn, pitch, grad, qa);
    pixdata* pd = (pixdata*)malloc(sizeof(pixdata) * (size_t)bw * bh);
    int npd = build_pixdata(g, pd);
    for (int by = 0; by < nby; by++)
        for (int bx = 0; bx < nbx; bx++) {
            float* hist = out + ((size_t)by * nbx + bx) * hsz;
            memset(hist, 0, sizeof(float) * hsz);
            int py = by * g->stride_h, px = bx * g->stride_w;
            for (int k = 0; k < npd; k++) {                     /* hog.cpp:900-1112 */
                const pixdata* p = &pd[k];
                size_t o = ((size_t)(py + p->i) * w + px + p->j) * 2;
                float a0 = grad[o], a1 = grad[o + 1];
                int h0 = qa[o], h1 = qa[o + 1];
                for (int c = 0; c < p->n; c++) {
                    float wgt = p->gw * p->hw[c];
                    float* hh = hist + p->ofs[c];
                    float t0 = hh[h0] + a0 * wgt;
                    float t1 = hh[h1] + a1 * wgt;
                    hh[h0] = t0; hh[h1] = t1;
                }
            }
            normalize_block(hist, hsz, (float)g->l2hys);
        }
    free(pd); free(qa); free(grad);
    return (long)nbx * nby;
}

/* ------------------------------------------------------------------ descriptors / scores */
static int parts_within(int size, int part, int stride) { return size >= part ? (size - part) / stride + 1 : 0; }

long hogo_compute(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                  int win_stride_x, int win_stride_y, float* descriptors)
{
    int wx = parts_within(w, g->win_w, win_stride_x), wy = parts_within(h, g->win_h, win_stride_y);
    if (wx <= 0 || wy <= 0) return 0;
    const int hsz = (g->block_w / g->cell_w) * (g->block_h / g->cell_h) * g->nbins;
    const int nbw = (g->win_w - g->block_w) / g->stride_w + 1, nbh = (g->win_h - g->block_h) / g->stride_h + 1;
    int nbx, nby;
    float* hists = (float*)malloc(sizeof(float) * hsz * (size_t)parts_within(w, g->block_w, g->stride_w)
                                  * (size_t)parts_within(h, g->block_h, g->stride_h));
    hogo_block_hists(g, img, w, h, cn, pitch, hists, &nbx, &nby);
    const size_t dsize = (size_t)nbw * nbh * hsz;
    for (int y = 0; y < wy; y++)
        for (int x = 0; x < wx; x++) {
            float* d = descriptors + ((size_t)y * wx + x) * dsize;
            int bx0 = x * win_stride_x / g->stride_w, by0 = y * win_stride_y / g->stride_h;
            for (int j = 0; j < nbw; j++)           /* blockData order: hog.cpp:851-857 (column-major) */
                for (int i = 0; i < nbh; i++)
                    memcpy(d + ((size_t)j * nbh + i) * hsz,
                           hists + ((size_t)(by0 + i) * nbx + bx0 + j) * hsz, sizeof(float) * hsz);
        }
    free(hists);
    return (long)wx * wy;
}

long hogo_scores(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                 const float* svm, int nsvm, int win_stride_x, int win_stride_y,
                 double* scores, int* wins_x, int* wins_y)
{
    int wx = parts_within(w, g->win_w, win_stride_x), wy = parts_within(h, g->win_h, win_stride_y);
    if (wx <= 0 || wy <= 0) wx = wy = 0;
    if (wins_x) *wins_x = wx;
    if (wins_y) *wins_y = wy;
    if (!wx) return 0;
    const int hsz = (g->block_w / g->cell_w) * (g->block_h / g->cell_h) * g->nbins;
    const int nbw = (g->win_w - g->block_w) / g->stride_w + 1, nbh = (g->win_h - g->block_h) / g->stride_h + 1;
    const int dsize = nbw * nbh * hsz;
    const double rho = nsvm > dsize ? svm[dsize] : 0;            /* hog.cpp:1688 */
    int nbx, nby;
    float* hists = (float*)malloc(sizeof(float) * hsz * (size_t)parts_within(w, g->block_w, g->stride_w)
                                  * (size_t)parts_within(h, g->block_h, g->stride_h));
    hogo_block_hists(g, img, w, h, cn, pitch, hists, &nbx, &nby);
    for (int y = 0; y < wy; y++)
        for (int x = 0; x < wx; x++) {
            int bx0 = x * win_stride_x / g->stride_w, by0 = y * win_stride_y / g->stride_h;
            double s = rho;
            const float* sv = svm;
            for (int j = 0; j < nbw; j++)
                for (int i = 0; i < nbh; i++, sv += hsz) {       /* hog.cpp:1713-1760 (SSE2 lanes) */
                    const float* vec = hists + ((size_t)(by0 + i) * nbx + bx0 + j) * hsz;
                    float part[4];
                    int k;
                    for (int l = 0; l < 4; l++) part[l] = sv[l] * vec[l];
                    for (k = 4; k <= hsz - 4; k += 4)
                        for (int l = 0; l < 4; l++) part[l] = part[l] + vec[k + l] * sv[k + l];
                    double t0 = part[0] + part[1];
                    double t1 = part[2] + part[3];
                    s += t0 + t1;
                    for (; k < hsz; k++) s += vec[k] * sv[k];
                }
            scores[(size_t)y * wx + x] = s;
        }
    free(hists);
    return (long)wx * wy;
}

/* ------------------------------------------------------------------ groupRectangles */
static int similar_rects(const int* a, const int* b, double eps)    /* objdetect.hpp:125-132 */
{
    double delta = eps * (imin(a[2], b[2]) + imin(a[3], b[3])) * 0.5;
    return abs(a[0] - b[0]) <= delta && abs(a[1] - b[1]) <= delta &&
           abs(a[0] + a[2] - b[0] - b[2]) <= delta && abs(a[1] + a[3] - b[1] - b[3]) <= delta;
}

long hogo_group_rectangles(int* rects, long n, int group_threshold, double eps)
{
    if (group_threshold <= 0 || n == 0) return n;                   /* cascadedetect.cpp:69-79 */
    int N = (int)n;
    /* cv::partition, operations.hpp:481-569 */
    int* parent = (int*)malloc(sizeof(int) * (size_t)N), *rank = (int*)malloc(sizeof(int) * (size_t)N);
    int* labels = (int*)malloc(sizeof(int) * (size_t)N);
    for (int i = 0; i < N; i++) { parent[i] = -1; rank[i] = 0; }
    for (int i = 0; i < N; i++) {
        int root = i;
        while (parent[root] >= 0) root = parent[root];
        for (int j = 0; j < N; j++) {
            if (i == j || !similar_rects(rects + 4 * i, rects + 4 * j, eps)) continue;
            int root2 = j;
            while (parent[root2] >= 0) root2 = parent[root2];
            if (root2 != root) {
                int r1 = rank[root], r2 = rank[root2];
                if (r1 > r2) parent[root2] = root;
                else { parent[root] = root2; rank[root2] += r1 == r2; root = root2; }
                int k = j, p;
                while ((p = parent[k]) >= 0) { parent[k] = root; k = p; }
                k = i;
                while ((p = parent[k]) >= 0) { parent[k] = root; k = p; }
            }
        }
    }
    int nclasses = 0;
    for (int i = 0; i < N; i++) {
        int root = i;
        while (parent[root] >= 0) root = parent[root];
        if (rank[root] >= 0) rank[root] = ~nclasses++;
        labels[i] = ~rank[root];
    }
    /* cascadedetect.cpp:83-174 */
    int* sum = (int*)calloc((size_t)nclasses * 4, sizeof(int));
    int* cnt = (int*)calloc((size_t)nclasses, sizeof(int));
    for (int i = 0; i < N; i++) {
        int c = labels[i];
        for (int k = 0; k < 4; k++) sum[4 * c + k] += rects[4 * i + k];
        cnt[c]++;
    }
    for (int c = 0; c < nclasses; c++) {
        float s = 1.f / cnt[c];
        for (int k = 0; k < 4; k++) sum[4 * c + k] = cv_roundf(sum[4 * c + k] * s);
    }
    long m = 0;
    for (int i = 0; i < nclasses; i++) {
        const int* r1 = sum + 4 * i;
        int n1 = cnt[i], j;
        if (n1 <= group_threshold) continue;
        for (j = 0; j < nclasses; j++) {
            int n2 = cnt[j];
            if (j == i || n2 <= group_threshold) continue;
            const int* r2 = sum + 4 * j;
            int dx = cv_round(r2[2] * eps), dy = cv_round(r2[3] * eps);
            if (r1[0] >= r2[0] - dx && r1[1] >= r2[1] - dy &&
                r1[0] + r1[2] <= r2[0] + r2[2] + dx && r1[1] + r1[3] <= r2[1] + r2[3] + dy &&
                (n2 > imax(3, n1) || n1 < 3))
                break;
        }
        if (j == nclasses) { memcpy(rects + 4 * m, r1, sizeof(int) * 4); m++; }
    }
    free(sum); free(cnt); free(parent); free(rank); free(labels);
    return m;
}

/* ------------------------------------------------------------------ whole path */
long hogo_detect_multiscale(const hogo_geom* g, const unsigned char* img, int w, int h, int cn, long pitch,
                            const float* svm, int nsvm, double hit_threshold,
                            int win_stride_x, int win_stride_y, double scale0, int nlevels,
                            int group_threshold, int* rects, double* conf, long capacity)
{
    enum { MAXLV = 1024 };
    double scales[MAXLV]; int lw[MAXLV], lh[MAXLV];
    int nl = hogo_levels(w, h, g->win_w, g->win_h, scale0, nlevels, scales, lw, lh, MAXLV);
    long n = 0;
    unsigned char* lvl = (unsigned char*)malloc((size_t)w * h * cn);
    double* sc = (double*)malloc(sizeof(double) * ((size_t)w * h / 64 + 16));
    for (int l = 0; l < nl; l++) {
        const unsigned char* p = img; long pp = pitch;
        if (lw[l] != w || lh[l] != h) {                            /* cudaobjdetect hog.cpp:453-468 */
            hogo_resize_linear(img, w, h, cn, pitch, lvl, lw[l], lh[l], (long)lw[l] * cn);
            p = lvl; pp = (long)lw[l] * cn;
        }
        int wx, wy;
        hogo_scores(g, p, lw[l], lh[l], cn, pp, svm, nsvm, win_stride_x, win_stride_y, sc, &wx, &wy);
        double s = scales[l];
        int sw_ = cv_round(g->win_w * s), sh_ = cv_round(g->win_h * s);   /* :471-472 */
        for (int y = 0; y < wy; y++)
            for (int x = 0; x < wx; x++) {
                double v = sc[(size_t)y * wx + x];
                if (!(v >= hit_threshold)) continue;
                if (n >= capacity) { free(lvl); free(sc); return -1; }
                rects[4 * n] = cv_round(x * win_stride_x * s);     /* Point2d(hit)*scale -> Point: cvRound */
                rects[4 * n + 1] = cv_round(y * win_stride_y * s);
                rects[4 * n + 2] = sw_; rects[4 * n + 3] = sh_;
                if (conf) conf[n] = v;
                n++;
            }
    }
    free(lvl); free(sc);
    if (group_threshold > 0) n = hogo_group_rectangles(rects, n, group_threshold, 0.2);   /* :484-487 */
    return n;
}
