/*
 * oracle/synth_oracle.c -- TEST / BENCH INFRASTRUCTURE ONLY: a HOST generator of the synthetic scan-pair workload of
 * SURVEY 8d (C4 / C5), so that bench.py's reference arm (`--impl reference`) builds its input without loading the
 * product library.  It is the twin of the device generator (laser_slam_b200/csrc/synth_sm100.cu): the same world model
 * and the same distributions -- an axis-aligned room U(6,20) x U(6,20) m with 4..12 boxes of 0.3..2 m, pose A uniform in
 * free space (0.5 m clearance), pose B = A (+) U(+-dmax_xy, +-dmax_xy, +-dmax_th), both scans ray cast, N(0, noise)
 * range noise, 1 mm quantisation, dropouts, everything beyond max range invalid (= max_range + 1 cm, as readCarmon
 * marks it) -- with its own counter-based generator (splitmix64 of seed / pair / beam), so the pairs are statistically
 * the same workload but not bit-identical to the device generator's (that one draws from cuRAND's Philox).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { uint64_t s; } rng_t;
static uint64_t sm64(uint64_t *s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static rng_t rng_at(uint64_t seed, uint64_t a, uint64_t b) {
    rng_t r;
    uint64_t s = seed ^ (a * 0xD1342543DE82EF95ull) ^ (b * 0xA0761D6478BD642Full);
    sm64(&s);
    r.s = s;
    return r;
}
static float uni(rng_t *r) { return (float)((sm64(&r->s) >> 40) + 1) * (1.0f / 16777216.0f); } /* (0, 1] */
static float nrm(rng_t *r) {
    float u1 = uni(r), u2 = uni(r);
    return sqrtf(-2.0f * logf(u1)) * cosf(6.2831853f * u2);
}

typedef struct {
    float W, H;
    int nbox;
    float bx0[12], by0[12], bx1[12], by1[12];
    float ax, ay, ath, bx, by, bth;
} world_t;

static int is_free(const world_t *w, float x, float y, float clr) {
    if (x < clr || y < clr || x > w->W - clr || y > w->H - clr) return 0;
    for (int b = 0; b < w->nbox; b++)
        if (x > w->bx0[b] - clr && x < w->bx1[b] + clr && y > w->by0[b] - clr && y < w->by1[b] + clr) return 0;
    return 1;
}

static float cast(const world_t *w, float ox, float oy, float dx, float dy) {
    float tx = dx > 0 ? (w->W - ox) / dx : (dx < 0 ? (0.f - ox) / dx : 1e30f);
    float ty = dy > 0 ? (w->H - oy) / dy : (dy < 0 ? (0.f - oy) / dy : 1e30f);
    float t = fminf(tx, ty);
    for (int b = 0; b < w->nbox; b++) {
        float t0 = 0.f, t1 = t;
        int hit = 1;
        if (dx != 0.f) {
            float a = (w->bx0[b] - ox) / dx, c = (w->bx1[b] - ox) / dx;
            t0 = fmaxf(t0, fminf(a, c)); t1 = fminf(t1, fmaxf(a, c));
        } else if (ox < w->bx0[b] || ox > w->bx1[b]) hit = 0;
        if (dy != 0.f) {
            float a = (w->by0[b] - oy) / dy, c = (w->by1[b] - oy) / dy;
            t0 = fmaxf(t0, fminf(a, c)); t1 = fminf(t1, fmaxf(a, c));
        } else if (oy < w->by0[b] || oy > w->by1[b]) hit = 0;
        if (hit && t0 <= t1 && t0 > 0.f) t = fminf(t, t0);
    }
    return t;
}

typedef struct {
    float *scans; double *truth;
    long long first_pair; int npairs, n;
    float fi_min, dfi; int max_range_cm;
    float dmax_xy, dmax_th, noise_m, dropout;
    uint64_t seed;
    int lo, hi;
} job_t;

static void one_pair(const job_t *j, int p) {
    world_t w;
    const uint64_t pid = (uint64_t)(j->first_pair + p);
    rng_t st = rng_at(j->seed, pid, 0);
    w.W = 6.f + 14.f * uni(&st);
    w.H = 6.f + 14.f * uni(&st);
    w.nbox = 4 + (int)(sm64(&st.s) % 9u);
    for (int b = 0; b < w.nbox; b++) {
        float sx = 0.3f + 1.7f * uni(&st), sy = 0.3f + 1.7f * uni(&st);
        float cx = uni(&st) * (w.W - sx), cy = uni(&st) * (w.H - sy);
        w.bx0[b] = cx; w.bx1[b] = cx + sx; w.by0[b] = cy; w.by1[b] = cy + sy;
    }
    float ddx = 0, ddy = 0, ddt = 0;
    for (int tries = 0; tries < 256; tries++) {
        w.ax = uni(&st) * w.W; w.ay = uni(&st) * w.H;
        w.ath = (uni(&st) * 2.f - 1.f) * 3.14159265f;
        ddx = (uni(&st) * 2.f - 1.f) * j->dmax_xy; ddy = (uni(&st) * 2.f - 1.f) * j->dmax_xy;
        ddt = (uni(&st) * 2.f - 1.f) * j->dmax_th;
        w.bx = w.ax + cosf(w.ath) * ddx - sinf(w.ath) * ddy;
        w.by = w.ay + sinf(w.ath) * ddx + cosf(w.ath) * ddy;
        w.bth = w.ath + ddt;
        if (is_free(&w, w.ax, w.ay, 0.5f) && is_free(&w, w.bx, w.by, 0.2f)) break;
        if (tries == 255) { w.nbox = 0; w.ax = w.W * 0.5f; w.ay = w.H * 0.5f; w.bx = w.ax + ddx; w.by = w.ay + ddy; w.ath = 0; w.bth = ddt; }
    }
    j->truth[3 * (size_t)p + 0] = ddx; j->truth[3 * (size_t)p + 1] = ddy; j->truth[3 * (size_t)p + 2] = ddt;
    for (int which = 0; which < 2; which++) {
        const float ox = which ? w.bx : w.ax, oy = which ? w.by : w.ay, oth = which ? w.bth : w.ath;
        float *row = j->scans + ((size_t)2 * p + which) * j->n;
        for (int i = 0; i < j->n; i++) {
            rng_t rs = rng_at(j->seed ^ 0x5851F42D4C957F2Dull, pid * 2 + which, (uint64_t)i + 1);
            float ang = oth + j->fi_min + j->dfi * (float)i;
            float t = cast(&w, ox, oy, cosf(ang), sinf(ang));
            float r = t + j->noise_m * nrm(&rs);
            r = rintf(r * 1000.f) * 0.001f;
            float cm = r * 100.f;
            int drop = uni(&rs) < j->dropout;
            if (drop || cm >= (float)j->max_range_cm || cm < 10.f) cm = (float)(j->max_range_cm + 1);
            row[i] = cm;
        }
    }
}

static void *worker(void *arg) {
    const job_t *j = (const job_t *)arg;
    for (int p = j->lo; p < j->hi; p++) one_pair(j, p);
    return 0;
}

/* scans [2*npairs][n] float cm (scans[2p] = ref, scans[2p+1] = sens), truth [npairs][3] = (dx, dy, dtheta) */
void orc_synth_pairs(float *scans, double *truth, long long first_pair, int npairs, int n, float fi_min, float dfi,
                     int max_range_cm, float dmax_xy_m, float dmax_th_rad, float noise_m, float dropout,
                     unsigned long long seed, int nthreads) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    if (nthreads > npairs) nthreads = npairs > 0 ? npairs : 1;
    pthread_t th[256];
    job_t jobs[256];
    for (int t = 0; t < nthreads; t++) {
        job_t j = {scans, truth, first_pair, npairs, n, fi_min, dfi, max_range_cm, dmax_xy_m, dmax_th_rad, noise_m, dropout,
                   (uint64_t)seed, (int)((long long)npairs * t / nthreads), (int)((long long)npairs * (t + 1) / nthreads)};
        jobs[t] = j;
        pthread_create(&th[t], 0, worker, &jobs[t]);
    }
    for (int t = 0; t < nthreads; t++) pthread_join(th[t], 0);
}
