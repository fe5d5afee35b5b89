/* TEST INFRASTRUCTURE — see gsdf_oracle.h.  Plain-C restatement of the reference hot path.
 * Every arithmetic step that feeds a branch or a round() keeps the reference's fp32 association order
 * (Eigen fixed-size 3-term reductions are right-nested: a0*b0 + (a1*b1 + a2*b2); no FMA). */
#include "gsdf_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ voxel hash (stands in for phmap) */
typedef struct {
    int32_t k[3];
    int32_t used;
    float dist, g[3], weight; /* SdfVoxel, sdf_voxel/SdfVoxel.h:45-57 */
} orc_slot;

struct orc_ctx {
    int variant, w, h, win;
    float K[9];
    float vs, vs_inv, T, inv_T; /* MapGradPixelSdf.h:99-103, Sdf.h:103-107 */
    float z_min, z_max;         /* Sdf.h:67-68 */
    float* planes;              /* 11 cached planes */
    orc_slot* tab;
    size_t cap, count;
    int world, rank; /* multi-GPU sharding protocol emulation: this context owns blocks with owner(key) == rank */
};

static uint64_t mix64(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
/* shard owner of the 8^3 block a voxel lies in: same function as block_owner() in the CUDA library
 * (gradient-sdf_b200/csrc/gsdf_device.cuh): packed 21-bit block coordinates, mix64, bits 40.. modulo world */
static int voxel_owner(const int32_t* v, int world) {
    const uint64_t bias = 1u << 20;
    const uint64_t key = (((uint64_t)(uint32_t)((v[0] >> 3) + (int32_t)bias)) & 0x1FFFFFull) |
                         ((((uint64_t)(uint32_t)((v[1] >> 3) + (int32_t)bias)) & 0x1FFFFFull) << 21) |
                         ((((uint64_t)(uint32_t)((v[2] >> 3) + (int32_t)bias)) & 0x1FFFFFull) << 42);
    return (int)((mix64(key) >> 40) % (uint64_t)world);
}

/* Which voxels a shard keeps (bit 0) and which of them it answers for (bit 1).  grad-sdf: the voxels of its own blocks.
 * base-sdf: a point's trilinear cell is evaluated on the shard that owns the block of its FLOOR corner, so that shard
 * also keeps copies of the low-face voxels (local coordinate 0) of the up-to-7 neighbour blocks at +1 (the halo);
 * every shard that keeps a voxel receives every sample of it, so the copies hold the same sums. */
static int shard_keeps(int variant, int world, int rank, const int32_t* v) {
    if (world <= 1 || voxel_owner(v, world) == rank) return 3;
    if (variant == 0) return 0;
    for (int o = 1; o < 8; ++o) {
        if (((o & 1) && (v[0] & 7)) || ((o & 2) && (v[1] & 7)) || ((o & 4) && (v[2] & 7))) continue;
        const int32_t nb[3] = {v[0] - 8 * (o & 1), v[1] - 8 * ((o >> 1) & 1), v[2] - 8 * (o >> 2)};
        if (voxel_owner(nb, world) == rank) return 1;
    }
    return 0;
}

static size_t key_hash(const int32_t* k) {
    /* the reference's hash (hash_map.h:44-52) is not observable; any hash gives the same key set */
    uint64_t a = (uint32_t)k[0], b = (uint32_t)k[1], c = (uint32_t)k[2];
    return (size_t)mix64(a * 0x9E3779B97F4A7C15ULL ^ mix64(b + 0x632BE59BD9B4E019ULL) ^ mix64(c * 0xD6E8FEB86659FD93ULL + 1));
}
static orc_slot* tab_find(const orc_ctx* c, const int32_t* k) {
    size_t i = key_hash(k) & (c->cap - 1);
    for (;;) {
        orc_slot* s = &c->tab[i];
        if (!s->used) return NULL;
        if (s->k[0] == k[0] && s->k[1] == k[1] && s->k[2] == k[2]) return s;
        i = (i + 1) & (c->cap - 1);
    }
}
static void tab_grow(orc_ctx* c) {
    orc_slot* old = c->tab;
    size_t oc = c->cap;
    c->cap = oc * 2;
    c->tab = (orc_slot*)calloc(c->cap, sizeof(orc_slot));
    for (size_t j = 0; j < oc; ++j) {
        if (!old[j].used) continue;
        size_t i = key_hash(old[j].k) & (c->cap - 1);
        while (c->tab[i].used) i = (i + 1) & (c->cap - 1);
        c->tab[i] = old[j];
    }
    free(old);
}
/* operator[] : find or zero-initialised insert */
static orc_slot* tab_get(orc_ctx* c, const int32_t* k) {
    if ((c->count + 1) * 10 > c->cap * 6) tab_grow(c);
    size_t i = key_hash(k) & (c->cap - 1);
    for (;;) {
        orc_slot* s = &c->tab[i];
        if (!s->used) {
            memset(s, 0, sizeof(*s));
            s->used = 1; s->k[0] = k[0]; s->k[1] = k[1]; s->k[2] = k[2];
            c->count++;
            return s;
        }
        if (s->k[0] == k[0] && s->k[1] == k[1] && s->k[2] == k[2]) return s;
        i = (i + 1) & (c->cap - 1);
    }
}

/* ------------------------------------------------------------------ small fp32 helpers */
static float dot3(const float* a, const float* b) { /* Eigen order, Appendix A of SURVEY.md */
    const float p0 = a[0] * b[0], p1 = a[1] * b[1], p2 = a[2] * b[2];
    const float s = p1 + p2;
    return p0 + s;
}
static void matvec3(const float* R, const float* v, float* o) { /* row-major R */
    o[0] = dot3(R, v); o[1] = dot3(R + 3, v); o[2] = dot3(R + 6, v);
}
static void float2vox(const orc_ctx* c, const float* p, int32_t* vi) { /* MapGradPixelSdf.h:74-77 */
    for (int i = 0; i < 3; ++i) { const float pv = c->vs_inv * p[i]; vi[i] = (int32_t)roundf(pv); }
}
static void vox2float(const orc_ctx* c, const int32_t* vi, float* p) { /* MapGradPixelSdf.h:79-81 */
    for (int i = 0; i < 3; ++i) p[i] = c->vs * (float)vi[i];
}
static float sdf_truncate(const orc_ctx* c, float sdf) { /* Sdf.h:72-74 */
    const float m = c->T < sdf ? c->T : sdf; /* std::min(T_, sdf) */
    return -c->T < m ? m : -c->T;            /* std::max(-T_, m) */
}
static float sdf_weight(const orc_ctx* c, float sdf) { /* Sdf.h:76-85 */
    float w = 0.f;
    if (sdf <= 0.) w = 1.f;
    else if (sdf <= c->T) w = 1.f - sdf * c->inv_T;
    return w;
}

static int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) { if (p < 0) p = -p; else p = 2 * (len - 1) - p; }
    return p;
}

/* unnormalised box sum, REFLECT_101, double accumulation: rows left->right, then columns top->bottom
 * (OpenCV boxFilter semantics as fixed by oracle/shim/opencv2/imgproc/imgproc.hpp) */
static void box_sum_d(const double* src, double* dst, int w, int h, int win) {
    const int r = win / 2;
    double* rows = (double*)malloc(sizeof(double) * (size_t)w * h);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            double s = 0.0;
            for (int d = -r; d <= r; ++d) s += src[(size_t)y * w + reflect101(x + d, w)];
            rows[(size_t)y * w + x] = s;
        }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            double s = 0.0;
            for (int d = -r; d <= r; ++d) s += rows[(size_t)reflect101(y + d, h) * w + x];
            dst[(size_t)y * w + x] = s;
        }
    free(rows);
}

/* ------------------------------------------------------------------ NormalEstimator::cache */
static void build_cache(orc_ctx* c) { /* normals/NormalEstimator.h:81-154, all in double then cast */
    const int w = c->w, h = c->h;
    const size_t n = (size_t)w * h;
    const double fx_inv = 1. / (double)c->K[0], fy_inv = 1. / (double)c->K[4];
    const double cx = (double)c->K[2], cy = (double)c->K[5];
    double* buf = (double*)malloc(sizeof(double) * n * 17);
    double *x0 = buf, *y0 = buf + n, *nsi = buf + 2 * n, *x0n = buf + 3 * n, *y0n = buf + 4 * n;
    double *a11 = buf + 5 * n, *a12 = buf + 6 * n, *a22 = buf + 7 * n;
    double *M11 = buf + 8 * n, *M12 = buf + 9 * n, *M13 = buf + 10 * n, *M22 = buf + 11 * n, *M23 = buf + 12 * n,
           *M33 = buf + 13 * n;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const size_t i = (size_t)y * w + x;
            const double u = fx_inv * ((double)x - cx), v = fy_inv * ((double)y - cy); /* :96-100 */
            const double x_sq = u * u, y_sq = v * v, xy = u * v;
            const double n_sq = (1. + x_sq) + y_sq; /* :103 */
            const double inv = 1. / n_sq;           /* :104 */
            x0[i] = u; y0[i] = v; nsi[i] = inv;
            x0n[i] = u * inv; y0n[i] = v * inv;      /* :105-106 */
            a11[i] = x_sq * inv; a12[i] = xy * inv; a22[i] = y_sq * inv; /* :108-111 */
        }
    box_sum_d(a11, M11, w, h, c->win); box_sum_d(a12, M12, w, h, c->win); box_sum_d(x0n, M13, w, h, c->win);
    box_sum_d(a22, M22, w, h, c->win); box_sum_d(y0n, M23, w, h, c->win); box_sum_d(nsi, M33, w, h, c->win);
    float* P = c->planes;
    for (size_t i = 0; i < n; ++i) {
        /* :115-124 */
        const double det = (M11[i] * (M22[i] * M33[i]) + 2 * M12[i] * (M23[i] * M13[i])) -
                           ((M13[i] * (M13[i] * M22[i]) + M12[i] * (M12[i] * M33[i])) + M23[i] * (M23[i] * M11[i]));
        const double di = 1. / det;
        P[0 * n + i] = (float)x0[i]; P[1 * n + i] = (float)y0[i];
        P[2 * n + i] = (float)x0n[i]; P[3 * n + i] = (float)y0n[i]; P[4 * n + i] = (float)nsi[i];
        P[5 * n + i] = (float)(di * (M22[i] * M33[i] - M23[i] * M23[i]));  /* Q11 */
        P[6 * n + i] = (float)(di * (M13[i] * M23[i] - M12[i] * M33[i]));  /* Q12 */
        P[7 * n + i] = (float)(di * (M12[i] * M23[i] - M13[i] * M22[i]));  /* Q13 */
        P[8 * n + i] = (float)(di * (M11[i] * M33[i] - M13[i] * M13[i]));  /* Q22 */
        P[9 * n + i] = (float)(di * (M12[i] * M13[i] - M11[i] * M23[i]));  /* Q23 */
        P[10 * n + i] = (float)(di * (M11[i] * M22[i] - M12[i] * M12[i])); /* Q33 */
    }
    free(buf);
}

orc_ctx* orc_create(int variant, float voxel_size, float trunc_factor, int w, int h, const float* K9, int normal_win) {
    orc_ctx* c = (orc_ctx*)calloc(1, sizeof(orc_ctx));
    c->variant = variant; c->w = w; c->h = h; c->win = normal_win;
    memcpy(c->K, K9, sizeof(float) * 9);
    c->vs = voxel_size;
    c->vs_inv = (float)(1. / (double)voxel_size);      /* MapGradPixelSdf.h:102 */
    c->T = trunc_factor * voxel_size;                   /* main_scan_3d.cpp:191 */
    c->inv_T = (float)(1. / (double)c->T);              /* Sdf.h:105 */
    c->z_min = 0.5f; c->z_max = 3.5f;
    c->world = 1; c->rank = 0;
    c->planes = (float*)malloc(sizeof(float) * (size_t)w * h * 11);
    c->cap = 1u << 16; c->count = 0;
    c->tab = (orc_slot*)calloc(c->cap, sizeof(orc_slot));
    build_cache(c);
    return c;
}

void orc_destroy(orc_ctx* c) {
    if (!c) return;
    free(c->planes); free(c->tab); free(c);
}

void orc_set_zrange(orc_ctx* c, float z_min, float z_max) { c->z_min = z_min; c->z_max = z_max; }
void orc_set_shard(orc_ctx* c, int world, int rank) { c->world = world; c->rank = rank; }
int orc_voxel_owner(const int32_t* v3, int world) { return voxel_owner(v3, world); }

void orc_normal_cache(const orc_ctx* c, float* out11) { memcpy(out11, c->planes, sizeof(float) * (size_t)c->w * c->h * 11); }

/* ------------------------------------------------------------------ NormalEstimator::compute */
void orc_normals(const orc_ctx* c, const float* depth, float* nx, float* ny, float* nz) {
    const int w = c->w, h = c->h, r = c->win / 2;
    const size_t n = (size_t)w * h;
    const float *x0n = c->planes + 2 * n, *y0n = c->planes + 3 * n, *nsi = c->planes + 4 * n;
    const float *Q11 = c->planes + 5 * n, *Q12 = c->planes + 6 * n, *Q13 = c->planes + 7 * n, *Q22 = c->planes + 8 * n,
                *Q23 = c->planes + 9 * n, *Q33 = c->planes + 10 * n;
    double* rs = (double*)malloc(sizeof(double) * n * 3);
    float* a = (float*)malloc(sizeof(float) * n * 3);
    for (size_t i = 0; i < n; ++i) {
        const float zi = depth[i] != 0.f ? 1.f / depth[i] : 0.f; /* :181-187 */
        a[i] = x0n[i] * zi; a[n + i] = y0n[i] * zi; a[2 * n + i] = nsi[i] * zi; /* :191-193 (float products) */
    }
    for (int p = 0; p < 3; ++p)
        for (int y = 0; y < h; ++y)
            for (int x = 0; x < w; ++x) {
                double s = 0.0;
                for (int d = -r; d <= r; ++d) s += (double)a[p * n + (size_t)y * w + reflect101(x + d, w)];
                rs[p * n + (size_t)y * w + x] = s;
            }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const size_t i = (size_t)y * w + x;
            double s1 = 0.0, s2 = 0.0, s3 = 0.0;
            for (int d = -r; d <= r; ++d) {
                const size_t j = (size_t)reflect101(y + d, h) * w + x;
                s1 += rs[j]; s2 += rs[n + j]; s3 += rs[2 * n + j];
            }
            const float b1 = (float)s1, b2 = (float)s2, b3 = (float)s3;
            const float vx = (b1 * Q11[i] + b2 * Q12[i]) + b3 * Q13[i]; /* :195-197 */
            const float vy = (b1 * Q12[i] + b2 * Q22[i]) + b3 * Q23[i];
            const float vz = (b1 * Q13[i] + b2 * Q23[i]) + b3 * Q33[i];
            const float nn = sqrtf((vx * vx + vy * vy) + vz * vz);     /* :199 */
            nx[i] = vx / nn; ny[i] = vy / nn; nz[i] = vz / nn;          /* :201-203 */
        }
    free(rs); free(a);
}

/* ------------------------------------------------------------------ fusion */
long long orc_update(orc_ctx* c, const float* depth, const float* R, const float* t, long long* n_points) {
    const int w = c->w, h = c->h;
    const size_t n = (size_t)w * h;
    float* nrm = (float*)malloc(sizeof(float) * n * 3);
    orc_normals(c, depth, nrm, nrm + n, nrm + 2 * n); /* MapGradPixelSdfOmp.cpp:59 */
    const float *x_hom = c->planes, *y_hom = c->planes + n, *hom_inv = c->planes + 4 * n; /* :65-70 */
    const float Rt2[3] = {R[2], R[5], R[8]}; /* row 2 of R^T */
    const int factor = (int)floorf(c->T / c->vs); /* :78 */
    long long updates = 0, points = 0;
    for (size_t idx = 0; idx < n; ++idx) {
        const float z = depth[idx];
        if (z <= c->z_min || z >= c->z_max) continue; /* :89 */
        const float xy[3] = {x_hom[idx], y_hom[idx], 1.f};
        float Rxy[3], Rn[3];
        matvec3(R, xy, Rxy);
        const float nv[3] = {nrm[idx], nrm[n + idx], nrm[2 * n + idx]};
        matvec3(R, nv, Rn);
        if ((double)dot3(nv, nv) < .1) continue; /* :97 */
        const float d = dot3(nv, xy);
        if ((double)((d * d) * hom_inv[idx]) < .25) continue; /* :100 */
        for (float k = (float)-factor; k <= (float)factor; ++k) { /* :103 */
            const float s = z + k * c->vs;
            float p[3], q[3], vf[3];
            int32_t vi[3];
            for (int i = 0; i < 3; ++i) p[i] = s * Rxy[i] + t[i]; /* :105 */
            float2vox(c, p, vi);
            vox2float(c, vi, vf);
            for (int i = 0; i < 3; ++i) q[i] = vf[i] - t[i];
            const float sdf = dot3(Rt2, q) - z; /* :107-108 */
            const float wgt = sdf_weight(c, sdf);
            const int keep = wgt > 0 ? shard_keeps(c->variant, c->world, c->rank, vi) : 0;
            if (keep) {
                orc_slot* v = tab_get(c, vi); /* :114 */
                v->weight += wgt;
                v->dist += (sdf_truncate(c, sdf) - v->dist) * wgt / v->weight; /* :116 */
                if (c->variant == 0)
                    for (int i = 0; i < 3; ++i) v->g[i] += wgt * Rn[i]; /* :117 */
                updates += keep >> 1; /* halo copies are not counted twice */
            }
        }
        ++points;
    }
    free(nrm);
    if (n_points) *n_points = points;
    return updates;
}

/* sharded base-sdf: halo copies are neither counted nor downloaded (the union over shards is the unsharded map) */
static int slot_reported(const orc_ctx* c, const orc_slot* s) {
    return s->used && (c->world <= 1 || c->variant == 0 || voxel_owner(s->k, c->world) == c->rank);
}
long long orc_voxel_count(const orc_ctx* c) {
    if (c->world <= 1 || c->variant == 0) return (long long)c->count;
    long long n = 0;
    for (size_t j = 0; j < c->cap; ++j) n += slot_reported(c, &c->tab[j]);
    return n;
}

void orc_download(const orc_ctx* c, int32_t* keys, float* dist, float* weight, float* grad) {
    size_t i = 0;
    for (size_t j = 0; j < c->cap; ++j) {
        const orc_slot* s = &c->tab[j];
        if (!slot_reported(c, s)) continue;
        keys[3 * i] = s->k[0]; keys[3 * i + 1] = s->k[1]; keys[3 * i + 2] = s->k[2];
        dist[i] = s->dist; weight[i] = s->weight;
        grad[3 * i] = s->g[0]; grad[3 * i + 1] = s->g[1]; grad[3 * i + 2] = s->g[2];
        ++i;
    }
}

/* ------------------------------------------------------------------ point queries */
static float grad_weights(const orc_ctx* c, const float* p) { /* MapGradPixelSdf.h:117-125 */
    int32_t vi[3];
    float2vox(c, p, vi);
    if (c->world > 1 && voxel_owner(vi, c->world) != c->rank) return 0.f; /* counted on the owning shard */
    const orc_slot* s = tab_find(c, vi);
    return s ? s->weight : 0.f;
}
static float grad_tsdf(const orc_ctx* c, const float* p, float* grad) { /* MapGradPixelSdf.h:109-115 */
    int32_t vi[3];
    float vf[3], nrm[3], d[3], a[3];
    float2vox(c, p, vi);
    const orc_slot* s = tab_find(c, vi);
    const float sq = dot3(s->g, s->g);
    if (sq > 0.f) { const float l = sqrtf(sq); for (int i = 0; i < 3; ++i) nrm[i] = s->g[i] / l; } /* Dot.h:121-131 */
    else { for (int i = 0; i < 3; ++i) nrm[i] = s->g[i]; }
    vox2float(c, vi, vf);
    for (int i = 0; i < 3; ++i) { a[i] = 1.2f * nrm[i]; d[i] = vf[i] - p[i]; } /* Eigen casts the literal to float */
    if (grad) { grad[0] = a[0]; grad[1] = a[1]; grad[2] = a[2]; }
    /* `1.2*v.grad.normalized().dot(..)` parses as 1.2 * (float dot): a DOUBLE product and sum, rounded on return */
    return (float)((double)s->dist + 1.2 * (double)dot3(nrm, d));
}

static float base_weights(const orc_ctx* c, const float* p) { /* MapPixelSdf.h:118-143 */
    int32_t m[3], k[3];
    for (int i = 0; i < 3; ++i) m[i] = (int32_t)floorf(c->vs_inv * p[i]);
    if (c->world > 1 && voxel_owner(m, c->world) != c->rank) return 0.f; /* counted on the shard owning the floor corner */
    for (int dz = 0; dz < 2; ++dz) for (int dy = 0; dy < 2; ++dy) for (int dx = 0; dx < 2; ++dx) {
        k[0] = m[0] + dx; k[1] = m[1] + dy; k[2] = m[2] + dz;
        if (!tab_find(c, k)) return 0.f;
    }
    float2vox(c, p, k);
    return tab_find(c, k)->weight;
}
static float base_interp3(const orc_ctx* c, const float* p, float extrap, float* grad) { /* MapPixelSdfOmp.cpp:44-113 */
    float pv[3], f[3], d[8];
    int32_t m[3], k[3];
    int found = 0;
    for (int i = 0; i < 3; ++i) { pv[i] = c->vs_inv * p[i]; m[i] = (int32_t)floorf(pv[i]); f[i] = pv[i] - (float)m[i]; }
    for (int k0 = 0; k0 < 2; ++k0) for (int j0 = 0; j0 < 2; ++j0) for (int i0 = 0; i0 < 2; ++i0) {
        k[0] = m[0] + i0; k[1] = m[1] + j0; k[2] = m[2] + k0;
        const orc_slot* s = tab_find(c, k);
        d[i0 + j0 * 2 + k0 * 4] = s ? s->dist : extrap;
        found += s ? 1 : 0;
    }
    if (found == 0) return extrap; /* :71-72 */
    if (found < 8) return 0.0f;    /* :110-111 */
    const float dx = f[0], dy = f[1], dz = f[2];
    const float d01 = (1 - dx) * d[0] + dx * d[1], d23 = (1 - dx) * d[2] + dx * d[3];
    const float d45 = (1 - dx) * d[4] + dx * d[5], d67 = (1 - dx) * d[6] + dx * d[7];
    const float d02 = (1 - dy) * d[0] + dy * d[2], d13 = (1 - dy) * d[1] + dy * d[3];
    const float d46 = (1 - dy) * d[4] + dy * d[6], d57 = (1 - dy) * d[5] + dy * d[7];
    grad[0] = c->vs_inv * ((((1 - dz) * d13 + dz * d57) - (1 - dz) * d02) - dz * d46); /* :97 */
    grad[1] = c->vs_inv * ((((1 - dz) * d23 + dz * d67) - (1 - dz) * d01) - dz * d45); /* :98 */
    grad[2] = c->vs_inv * ((((1 - dy) * d45 + dy * d67) - (1 - dy) * d01) - dy * d23); /* :99 */
    const float dy0 = (1 - dy) * d01 + dy * d23, dy1 = (1 - dy) * d45 + dy * d67;
    return (1 - dz) * dy0 + dz * dy1; /* :107 */
}

static float q_weights(const orc_ctx* c, const float* p) { return c->variant == 0 ? grad_weights(c, p) : base_weights(c, p); }
static float q_tsdf(const orc_ctx* c, const float* p, float* grad) {
    if (c->variant == 0) return grad_tsdf(c, p, grad);
    grad[0] = grad[1] = grad[2] = 0.f; /* MapPixelSdf.h:108 */
    return base_interp3(c, p, -c->T, grad);
}

void orc_query(const orc_ctx* c, const float* pts, long long n, float* phi, float* grad, float* w) {
    for (long long i = 0; i < n; ++i) {
        const float* p = pts + 3 * i;
        const float w0 = q_weights(c, p);
        float g[3] = {0.f, 0.f, 0.f}, f = 0.f;
        if (w0 > 0) f = q_tsdf(c, p, g);
        w[i] = w0; phi[i] = f;
        grad[3 * i] = g[0]; grad[3 * i + 1] = g[1]; grad[3 * i + 2] = g[2];
    }
}

/* ------------------------------------------------------------------ tracker */
static void accumulate_full(const orc_ctx* c, const float* depth, const float* R, const float* t, int sampling,
                            float* g6, float* H36, float* E_out, float* cnt_out) {
    const int w = c->w, h = c->h;
    const float fx_inv = 1.f / c->K[0], fy_inv = 1.f / c->K[4], cx = c->K[2], cy = c->K[5]; /* :48-52 */
    float E = 0.f;
    size_t counter = 0;
    memset(g6, 0, sizeof(float) * 6); memset(H36, 0, sizeof(float) * 36);
    for (int y = 0; y < h; y += sampling)
        for (int x = 0; x < w; x += sampling) {
            const float z = depth[(size_t)y * w + x];
            if (z <= c->z_min || z >= c->z_max) continue; /* :72-73 */
            const float x0 = ((float)x - cx) * fx_inv, y0 = ((float)y - cy) * fy_inv; /* :75-76 */
            const float pc[3] = {x0 * z, y0 * z, z};
            float p[3], gr[3], J[6];
            matvec3(R, pc, p);
            for (int i = 0; i < 3; ++i) p[i] = p[i] + t[i]; /* :78 */
            const float w0 = q_weights(c, p);
            if (w0 > 0) {
                const float phi = q_tsdf(c, p, gr); /* :83 */
                E += phi * phi;
                J[0] = gr[0]; J[1] = gr[1]; J[2] = gr[2];
                J[3] = p[1] * gr[2] - p[2] * gr[1]; /* point.cross(grad), :86 */
                J[4] = p[2] * gr[0] - p[0] * gr[2];
                J[5] = p[0] * gr[1] - p[1] * gr[0];
                for (int i = 0; i < 6; ++i) g6[i] += phi * J[i]; /* :87 */
                for (int i = 0; i < 6; ++i)
                    for (int j = 0; j < 6; ++j) H36[6 * i + j] += J[i] * J[j]; /* :88 */
                ++counter;
            }
        }
    *E_out = E; *cnt_out = (float)counter;
}

void orc_accumulate(const orc_ctx* c, const float* depth, const float* R9, const float* t3, int sampling, float* out29) {
    float g[6], H[36], E, cnt;
    accumulate_full(c, depth, R9, t3, sampling, g, H, &E, &cnt);
    int o = 0;
    for (int i = 0; i < 6; ++i) out29[o++] = g[i];
    for (int i = 0; i < 6; ++i) for (int j = i; j < 6; ++j) out29[o++] = H[6 * i + j];
    out29[o++] = E; out29[o++] = cnt;
}

int orc_llt_solve(const float* H36, const float* g6, float* xi6) {
    /* Eigen LLT, unblocked lower Cholesky in place (stops at the first non-positive pivot, leaving the rest of
     * the matrix as is), then forward/back substitution (RigidPointOptimizerOmp.cpp:93) */
    float A[36], y[6];
    int ok = 1;
    memcpy(A, H36, sizeof(A));
    for (int k = 0; k < 6; ++k) {
        float x = A[6 * k + k];
        for (int j = 0; j < k; ++j) x -= A[6 * k + j] * A[6 * k + j];
        if (x <= 0.f) { ok = 0; break; }
        x = sqrtf(x);
        A[6 * k + k] = x;
        for (int i = k + 1; i < 6; ++i) {
            float s = A[6 * i + k];
            for (int j = 0; j < k; ++j) s -= A[6 * i + j] * A[6 * k + j];
            A[6 * i + k] = s / x;
        }
    }
    for (int i = 0; i < 6; ++i) {
        float s = g6[i];
        for (int j = 0; j < i; ++j) s -= A[6 * i + j] * y[j];
        y[i] = s / A[6 * i + i];
    }
    for (int i = 5; i >= 0; --i) {
        float s = y[i];
        for (int j = i + 1; j < 6; ++j) s -= A[6 * j + i] * xi6[j];
        xi6[i] = s / A[6 * i + i];
    }
    return ok;
}

/* ---- pose algebra: quaternion (x,y,z,w) + translation, as Sophus::SE3f stores it */
static void quat_normalize(float* q) { /* so3.hpp:297-303 */
    /* Eigen's 4-float squaredNorm is one SSE packet: horizontal add pairs lanes as (x^2+z^2)+(y^2+w^2) */
    const float len = sqrtf((q[0] * q[0] + q[2] * q[2]) + (q[1] * q[1] + q[3] * q[3]));
    for (int i = 0; i < 4; ++i) q[i] /= len;
}
static void quat_to_R(const float* q, float* R) { /* Eigen QuaternionBase::toRotationMatrix */
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    const float tx = 2.f * x, ty = 2.f * y, tz = 2.f * z;
    const float twx = tx * w, twy = ty * w, twz = tz * w;
    const float txx = tx * x, txy = ty * x, txz = tz * x, tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R[0] = 1.f - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
    R[3] = txy + twz; R[4] = 1.f - (txx + tzz); R[5] = tyz - twx;
    R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1.f - (txx + tyy);
}
static void quat_rotate(const float* q, const float* p, float* o) { /* so3.hpp:356-366 */
    const float* v = q;
    float uv[3] = {v[1] * p[2] - v[2] * p[1], v[2] * p[0] - v[0] * p[2], v[0] * p[1] - v[1] * p[0]};
    for (int i = 0; i < 3; ++i) uv[i] += uv[i];
    const float c2[3] = {v[1] * uv[2] - v[2] * uv[1], v[2] * uv[0] - v[0] * uv[2], v[0] * uv[1] - v[1] * uv[0]};
    for (int i = 0; i < 3; ++i) o[i] = (p[i] + q[3] * uv[i]) + c2[i];
}
static void quat_mul(const float* a, const float* b, float* o) { /* so3.hpp:325-339 (x,y,z,w storage) */
    const float ax = a[0], ay = a[1], az = a[2], aw = a[3], bx = b[0], by = b[1], bz = b[2], bw = b[3];
    o[3] = aw * bw - ax * bx - ay * by - az * bz;
    o[0] = aw * bx + ax * bw + ay * bz - az * by;
    o[1] = aw * by + ay * bw + az * bx - ax * bz;
    o[2] = aw * bz + az * bw + ax * by - ay * bx;
    quat_normalize(o); /* SO3(quaternion) constructor, so3.hpp:481-487 */
}
static void se3_exp(const float* a, float* q, float* tr) { /* se3.hpp:761-782, so3.hpp:583-619 */
    const float* om = a + 3;
    const float eps = 1e-5f;
    const float theta_sq = dot3(om, om);
    float theta, imag, real;
    if (theta_sq < eps * eps) {
        theta = 0.f;
        const float p4 = theta_sq * theta_sq;
        imag = 0.5f - (float)(1.0 / 48.0) * theta_sq + (float)(1.0 / 3840.0) * p4;
        real = 1.f - (float)(1.0 / 8.0) * theta_sq + (float)(1.0 / 384.0) * p4;
    } else {
        theta = sqrtf(theta_sq);
        const float half = 0.5f * theta;
        imag = sinf(half) / theta;
        real = cosf(half);
    }
    q[0] = imag * om[0]; q[1] = imag * om[1]; q[2] = imag * om[2]; q[3] = real;
    const float O[9] = {0.f, -om[2], om[1], om[2], 0.f, -om[0], -om[1], om[0], 0.f}; /* hat(omega) */
    float O2[9], V[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const float col[3] = {O[j], O[3 + j], O[6 + j]};
            O2[3 * i + j] = dot3(O + 3 * i, col);
        }
    if (theta < eps) {
        quat_to_R(q, V);
    } else {
        const float tsq = theta * theta;
        const float c1 = (1.f - cosf(theta)) / tsq, c2 = (theta - sinf(theta)) / (tsq * theta);
        for (int i = 0; i < 9; ++i) V[i] = ((i % 4 == 0 ? 1.f : 0.f) + c1 * O[i]) + c2 * O2[i];
    }
    matvec3(V, a, tr);
}

void orc_pose_matrix(const float* pose7, float* R9, float* pose7_normalised) {
    float q[4] = {pose7[3], pose7[4], pose7[5], pose7[6]};
    quat_normalize(q);
    quat_to_R(q, R9);
    if (pose7_normalised) {
        pose7_normalised[0] = pose7[0]; pose7_normalised[1] = pose7[1]; pose7_normalised[2] = pose7[2];
        for (int i = 0; i < 4; ++i) pose7_normalised[3 + i] = q[i];
    }
}

void orc_apply_increment(const float* pose7_in, const float* xi6, float* pose7_out) {
    /* pose = SE3::exp(-xi) * pose (RigidPointOptimizerOmp.cpp:102; product se3.hpp: q = qa*qb, t = ta + qa*tb) */
    float a[6], qa[4], ta[3], qb[4] = {pose7_in[3], pose7_in[4], pose7_in[5], pose7_in[6]}, qo[4], rt[3];
    for (int i = 0; i < 6; ++i) a[i] = -xi6[i];
    se3_exp(a, qa, ta);
    quat_mul(qa, qb, qo);
    quat_rotate(qa, pose7_in, rt);
    for (int i = 0; i < 3; ++i) pose7_out[i] = ta[i] + rt[i];
    for (int i = 0; i < 4; ++i) pose7_out[3 + i] = qo[i];
}

int orc_optimize(const orc_ctx* c, const float* depth, float* pose7, int iters, float conv_threshold, float damping,
                 int sampling, int* iters_used) {
    const float thr_sq = conv_threshold * conv_threshold; /* RigidOptimizer.h:72 */
    for (int k = 0; k < iters; ++k) {
        float R[9], g[6], H[36], E, cnt, xi[6], nrm = 0.f;
        float q[4] = {pose7[3], pose7[4], pose7[5], pose7[6]};
        quat_to_R(q, R); /* pose_.rotationMatrix(), :59 */
        accumulate_full(c, depth, R, pose7, sampling, g, H, &E, &cnt);
        orc_llt_solve(H, g, xi);
        int has_nan = 0;
        for (int i = 0; i < 6; ++i) { xi[i] = damping * xi[i]; if (isnan(xi[i])) has_nan = 1; }
        /* squaredNorm of a 6-vector; order is irrelevant for the comparison at 1e-6 */
        for (int i = 0; i < 6; ++i) nrm += xi[i] * xi[i];
        if (nrm < thr_sq) { if (iters_used) *iters_used = k; return 1; } /* :95-98 */
        if (!has_nan) { float out[7]; orc_apply_increment(pose7, xi, out); memcpy(pose7, out, sizeof(out)); } /* :101-102 */
    }
    if (iters_used) *iters_used = iters;
    return 0;
}
