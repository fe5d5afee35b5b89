// Marching cubes over the voxel-block hash (Sdf::extract_mesh: MapGradPixelSdfOmp.cpp:132-183, MapPixelSdfOmp.cpp:197-248
// -> mesh/LayeredMarchingCubesNoColor.cpp:354-719).  The reference sweeps the map's bounding box two z-layers at a time
// with a hash find per cell corner; here every allocated 8^3 block is one unit of work: a CTA stages the block and
// the low faces of its up-to-7 neighbours at +1 (9^3 voxels) in shared memory, and one thread per cell applies the
// reference's rules:
//   * a cell is meshed only if all 8 corners have weight != 0 (LayeredMarchingCubesNoColor.cpp:593-636);
//   * corner numbering / bit order of the case index as computeLutIndex (:593-636): bit 1 = (i+1, j+1, k), 2 = (i+1, j, k),
//     4 = (i, j, k), 8 = (i, j+1, k), 16..128 the same at k+1; a bit is set where dist > iso (0);
//   * edge points by interpolate() (:639-662): the 1e-7 shortcuts, mu = float quotient widened to double, clamped to
//     [0, 1], val0 + mu * (val1 - val0) in double, rounded to float; corner positions by voxelToWorld() (:712-716):
//     float(i) * vs - (-(float(min)) * vs) with i relative to the bounding-box minimum;
//   * triangles in the order of the case table, a triangle is dropped when two of its points coincide (:676-693);
//     vertices are not shared between triangles (addVertex, :696-702).
// Triangles are emitted with the key (cell index in the reference's sweep order z, y, x; triangle number in the cell),
// and the host sorts by that key, so that the PLY lists vertices and faces exactly in the reference's order.
#pragma once

#include <limits.h>

#include "gsdf_device.cuh"

namespace gsdf {

// The marching-cubes case table (Lorensen & Cline; the public-domain "polygonise" triangle table that the reference
// embeds too, mesh/LayeredMarchingCubesNoColor.cpp:96-352), nibble-packed: bits 60..63 = number of edge indices n
// (0, 3, .. 15), nibble i < n = the i-th edge index.  tests/test_oracle.py checks all 256 rows against the table inside
// the reference build.  The reference's edgeTable (:67-94) is not stored: bit e of it is set iff the two corners of
// edge e differ in the case index, and only the edges a case's triangles name are ever interpolated.
__device__ const unsigned long long kTriTable[256] = {
    0x0000000000000000ull, 0x3000000000000380ull, 0x3000000000000910ull, 0x6000000000189381ull,
    0x3000000000000a21ull, 0x6000000000a21380ull, 0x6000000000920a29ull, 0x900000089a8a2382ull,
    0x30000000000002b3ull, 0x60000000000b82b0ull, 0x6000000000b32091ull, 0x9000000b89b912b1ull,
    0x60000000003ab1a3ull, 0x9000000ab8a801a0ull, 0x90000009ab9b3093ull, 0x6000000000b8aa89ull,
    0x3000000000000874ull, 0x6000000000437034ull, 0x6000000000748910ull, 0x9000000137174914ull,
    0x6000000000748a21ull, 0x9000000a21403743ull, 0x9000000748209a29ull, 0xc0004973727929a2ull,
    0x60000000002b3748ull, 0x900000040242b74bull, 0x9000000b32748109ull, 0xc0001292b9b49b74ull,
    0x9000000487ab31a3ull, 0xc0004b7401b41ab1ull, 0xc00030bab9b09874ull, 0x9000000ab99b4b74ull,
    0x3000000000000459ull, 0x6000000000380459ull, 0x6000000000051450ull, 0x9000000513538458ull,
    0x6000000000459a21ull, 0x9000000594a21803ull, 0x9000000204245a25ull, 0xc0008434535235a2ull,
    0x6000000000b32459ull, 0x9000000594b802b0ull, 0x9000000b32510450ull, 0xc000584b82852512ull,
    0x900000045931ab3aull, 0xc000ab81a8180594ull, 0xc00030bab5b05045ull, 0x9000000b8aa85845ull,
    0x6000000000975879ull, 0x9000000375359039ull, 0x9000000751710870ull, 0x6000000000753351ull,
    0x900000021a759879ull, 0xc00037503505921aull, 0xc00025a758528208ull, 0x90000007533525a2ull,
    0x90000002b3987597ull, 0xc000b72029279759ull, 0xc000751871810b32ull, 0x900000051771b12bull,
    0xc000b3a31a758859ull, 0xf0aba010b7905075ull, 0xf07570805a30b0abull, 0x60000000005b75abull,
    0x300000000000056aull, 0x60000000006a5380ull, 0x60000000006a5109ull, 0x90000006a5891381ull,
    0x6000000000162561ull, 0x9000000803621561ull, 0x9000000620609569ull, 0xc000823625285895ull,
    0x600000000056ab32ull, 0x900000056a02b80bull, 0x90000006a5b32910ull, 0xc000b892b92916a5ull,
    0x9000000315356b36ull, 0xc0006b51505b0b80ull, 0xc0009505606306b3ull, 0x900000089bb96956ull,
    0x60000000008746a5ull, 0x9000000a56374034ull, 0x90000007486a5091ull, 0xc00049737179156aull,
    0x9000000874156216ull, 0xc000743403625521ull, 0xc000620560509748ull, 0xf962695923497937ull,
    0x900000056a4872b3ull, 0xc000b720242746a5ull, 0xc0006a5b32874910ull, 0xf6a54b7b492b9129ull,
    0xc0006b51535b3748ull, 0xfb404b7b016b5b15ull, 0xf74836b630560950ull, 0xc0009b7974b96956ull,
    0x6000000000a4694aull, 0x9000000380a946a4ull, 0x900000004606a10aull, 0xc000a16468618138ull,
    0x9000000462421941ull, 0xc000462942921803ull, 0x6000000000624420ull, 0x9000000624428238ull,
    0x900000032b46a94aull, 0xc0006a4a94b82280ull, 0xc000a164606102b3ull, 0xf1b8b12184a16146ull,
    0xc00036b319639469ull, 0xf14641916b0181b8ull, 0x90000004600636b3ull, 0x600000000086b846ull,
    0x9000000a98a876a7ull, 0xc000a76a907a0370ull, 0xc0000818717a176aull, 0x900000037117a76aull,
    0xc000768981861621ull, 0xf937390976192962ull, 0x9000000206607087ull, 0x6000000000276237ull,
    0xc00076898a86ab32ull, 0xf7a9a76790b72702ull, 0xfb32a767a1871081ull, 0xc00017616a71b12bull,
    0xf63136b619768698ull, 0x600000000076b190ull, 0xc00006b0b3607087ull, 0x30000000000006b7ull,
    0x3000000000000b67ull, 0x600000000067b803ull, 0x600000000067b910ull, 0x900000067b138918ull,
    0x60000000007b621aull, 0x90000007b6803a21ull, 0x90000007b69a2092ull, 0xc00089a38a3a27b6ull,
    0x6000000000726327ull, 0x9000000026067807ull, 0x9000000910732672ull, 0xc000678891681261ull,
    0x900000073171a67aull, 0xc000801781a7167aull, 0xc0007a69a0a70730ull, 0x90000009a88a7a67ull,
    0x600000000068b486ull, 0x9000000640603b63ull, 0x9000000109648b68ull, 0xc00063b139369649ull,
    0x90000001a28b6486ull, 0xc000640b60b03a21ull, 0xc0009a2920b648b4ull, 0xf36463b34923a39aull,
    0x9000000264248328ull, 0x6000000000264240ull, 0xc000834642432091ull, 0x9000000642241491ull,
    0xc0001a6648168318ull, 0x900000040660a01aull, 0xf39a9303a6834364ull, 0x60000000004a649aull,
    0x6000000000b67594ull, 0x900000067b594380ull, 0x9000000b67045105ull, 0xc00051345343867bull,
    0x9000000b6721a459ull, 0xc000594380a217b6ull, 0xc000204a24a45b67ull, 0xf67b25a523453843ull,
    0x9000000945267327ull, 0xc000786260680459ull, 0xc000045051673263ull, 0xf851584812786826ull,
    0xc00073167161a459ull, 0xf459078701671a61ull, 0xfa737a6a305a4a04ull, 0xc000a84a458a7a67ull,
    0x900000098b9b6596ull, 0xc000590650360b63ull, 0xc000b65510b508b0ull, 0x90000001355363b6ull,
    0xc00065b8b9b59a21ull, 0xfa21965690b603b0ull, 0xf52025a50865b58bull, 0xc00035a3a25363b6ull,
    0xc000283265825985ull, 0x9000000260069659ull, 0xf826283865081851ull, 0x6000000000612651ull,
    0xf698965683a61631ull, 0xc00006505960a01aull, 0x6000000000a65830ull, 0x300000000000065aull,
    0x6000000000b57a5bull, 0x900000003857ba5bull, 0x9000000091ba57b5ull, 0xc0001381897ba57aull,
    0x900000015717b21bull, 0xc000b27571721380ull, 0xc0007b2209729579ull, 0xf289823295b27257ull,
    0x9000000573532a52ull, 0xc00052a578258028ull, 0xc0002a37353a5109ull, 0xf25752a278129289ull,
    0x6000000000573531ull, 0x9000000571170780ull, 0x9000000735539309ull, 0x6000000000795789ull,
    0x90000008ba8a5485ull, 0xc00003bba50b5405ull, 0xc00054aba8a48910ull, 0xf41314943b54a4baull,
    0xc0008548b2582152ull, 0xfb151b2b543b0b40ull, 0xf58b8545b2950520ull, 0x60000000003b2549ull,
    0xc000483543253a52ull, 0x90000000244252a5ull, 0xf910854583a532a3ull, 0xc0002492914252a5ull,
    0x9000000153358548ull, 0x6000000000501540ull, 0xc000530509358548ull, 0x3000000000000549ull,
    0x9000000ba9b947b4ull, 0xc000ba97b9794380ull, 0xc000b470414b1ba1ull, 0xf4bab474a1843413ull,
    0xc000219b294b97b4ull, 0xf3801b2b197b9479ull, 0x900000004224b47bull, 0xc00042343824b47bull,
    0xc000947732972a92ull, 0xf70207872a4797a9ull, 0xfa040a1a472a3a73ull, 0x60000000004782a1ull,
    0x9000000317714194ull, 0xc000178180714194ull, 0x6000000000347304ull, 0x3000000000000784ull,
    0x60000000008ba8a9ull, 0x9000000a9bb93903ull, 0x9000000ba88a0a10ull, 0x6000000000a3ba13ull,
    0x90000008b99b1b21ull, 0xc0009b2921b93903ull, 0x6000000000b08b20ull, 0x3000000000000b23ull,
    0x900000098aa82832ull, 0x60000000002902a9ull, 0xc0008a1810a82832ull, 0x30000000000002a1ull,
    0x6000000000819831ull, 0x3000000000000190ull, 0x3000000000000830ull, 0x0000000000000000ull,
};

// corner c of a cell (numbering of computeLutIndex): offsets from the cell's base voxel (i, j, k)
__device__ __forceinline__ void mc_corner(int c, int& dx, int& dy, int& dz) {
    dx = (c == 0 || c == 1 || c == 4 || c == 5) ? 1 : 0;
    dy = (c == 0 || c == 3 || c == 4 || c == 7) ? 1 : 0;
    dz = c >> 2;
}
// the two corners of edge e: 0:0-1 1:1-2 2:2-3 3:3-0 4:4-5 5:5-6 6:6-7 7:7-4 8:0-4 9:1-5 10:2-6 11:3-7
__device__ __forceinline__ void mc_edge(int e, int& a, int& b) {
    if (e < 4) { a = e; b = (e + 1) & 3; }
    else if (e < 8) { a = e; b = 4 + ((e + 1) & 3); }
    else { a = e - 8; b = e - 4; }
}

struct MeshBox {
    int min[3], dim[3];  // bounding box of the voxel keys with weight > 0 (computeIsoSurface, :361-385)
};

// bounding box of the live voxels: box[0..2] = min, box[3..5] = max (atomicMin / atomicMax on ints)
template <int VARIANT>
__global__ void __launch_bounds__(256)
mesh_bbox_kernel(MapView m, int* __restrict__ box) {
    const unsigned int warps_per_block = blockDim.x >> 5, lane = threadIdx.x & 31;
    int lo[3] = {INT_MAX, INT_MAX, INT_MAX}, hi[3] = {INT_MIN, INT_MIN, INT_MIN};
    for (unsigned int slot = blockIdx.x * warps_per_block + (threadIdx.x >> 5); slot <= m.cap_mask;
         slot += gridDim.x * warps_per_block) {
        const unsigned long long key = m.keys[slot];
        if (key == kEmptyKey) continue;
        int bx, by, bz;
        unpack_block(key, bx, by, bz);
        for (int v = lane; v < kBlockVoxels; v += 32) {
            const size_t addr = (size_t)slot * kBlockVoxels + v;
            const float W = VARIANT == 0 ? m.accW[addr] : m.acc2[addr].x;
            if (W > 0.f) {
                const int x = (bx << 3) | (v & 7), y = (by << 3) | ((v >> 3) & 7), z = (bz << 3) | (v >> 6);
                lo[0] = min(lo[0], x); lo[1] = min(lo[1], y); lo[2] = min(lo[2], z);
                hi[0] = max(hi[0], x); hi[1] = max(hi[1], y); hi[2] = max(hi[2], z);
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        lo[c] = __reduce_min_sync(0xffffffffu, lo[c]);
        hi[c] = __reduce_max_sync(0xffffffffu, hi[c]);
        if (lane == 0 && lo[c] != INT_MAX) { atomicMin(box + c, lo[c]); atomicMax(box + 3 + c, hi[c]); }
    }
}

// interpolate(), LayeredMarchingCubesNoColor.cpp:639-662, iso value 0
__device__ __forceinline__ void mc_interpolate(float t0, float t1, const float* v0, const float* v1, float* out) {
    const float* pick = nullptr;
    if ((double)fabsf(__fsub_rn(0.f, t0)) < 1e-7) pick = v0;
    else if ((double)fabsf(__fsub_rn(0.f, t1)) < 1e-7) pick = v1;
    else if ((double)fabsf(__fsub_rn(t0, t1)) < 1e-7) pick = v0;
    if (pick) { out[0] = pick[0]; out[1] = pick[1]; out[2] = pick[2]; return; }
    double mu = (double)__fdiv_rn(__fsub_rn(0.f, t0), __fsub_rn(t1, t0));
    if (mu > 1.0) mu = 1.0;
    else if (mu < 0) mu = 0.0;
#pragma unroll
    for (int c = 0; c < 3; ++c)
        out[c] = (float)__dadd_rn((double)v0[c], __dmul_rn(mu, (double)__fsub_rn(v1[c], v0[c])));
}

// One CTA (256 threads) per allocated block.  EMIT = false: only counts the triangles (total[0]); EMIT = true: writes
// {key, 9 floats} per triangle at a cursor (total[0]) while the cursor stays below max_tris.
template <int VARIANT, bool EMIT>
__global__ void __launch_bounds__(256)
mesh_kernel(MapView m, MeshBox box, float vs, unsigned long long* __restrict__ total, long long max_tris,
            unsigned long long* __restrict__ keys, float* __restrict__ verts) {
    __shared__ float s_w[9 * 9 * 9], s_d[9 * 9 * 9];
    __shared__ int s_nslot[8];
    const int lane = threadIdx.x & 31;
    for (unsigned int slot = blockIdx.x; slot <= m.cap_mask; slot += gridDim.x) {
        const unsigned long long key = m.keys[slot];
        if (key == kEmptyKey) continue;  // uniform over the CTA
        int bx, by, bz;
        unpack_block(key, bx, by, bz);
        __syncthreads();  // the previous block's cells are done with the shared arrays
        if (threadIdx.x < 8) {
            const int o = threadIdx.x;
            s_nslot[o] = o == 0 ? (int)slot : map_find(m, pack_block(bx + (o & 1), by + ((o >> 1) & 1), bz + (o >> 2)));
        }
        __syncthreads();
        for (int i = threadIdx.x; i < 729; i += 256) {
            const int x = i % 9, y = (i / 9) % 9, z = i / 81;
            const int ns = s_nslot[(x >> 3) | ((y >> 3) << 1) | ((z >> 3) << 2)];
            float W = 0.f, S = 0.f;
            if (ns >= 0) {
                const size_t addr = (size_t)ns * kBlockVoxels + ((x & 7) | ((y & 7) << 3) | ((z & 7) << 6));
                if (VARIANT == 0) { W = m.accW[addr]; S = m.acc4[addr].x; }
                else { const float2 a = m.acc2[addr]; W = a.x; S = a.y; }
            }
            s_w[i] = W;
            s_d[i] = W > 0.f ? __fdiv_rn(S, W) : 0.f;  // dist as gsdf_download reports it
        }
        __syncthreads();
        for (int cell = threadIdx.x; cell < 512; cell += 256) {
            const int lx = cell & 7, ly = (cell >> 3) & 7, lz = cell >> 6;
            float d[8];
            bool all = true;
            int cube = 0;
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                int dx, dy, dz;
                mc_corner(c, dx, dy, dz);
                const int i = (lx + dx) + 9 * (ly + dy) + 81 * (lz + dz);
                all = all && (s_w[i] != 0.f);
                d[c] = s_d[i];
                cube |= d[c] > 0.f ? (1 << c) : 0;
            }
            if (!all) cube = 0;
            float tri[5][9];
            int ntri = 0, tpos[5];
            if (cube != 0 && cube != 255) {
                // the cell's base voxel relative to the bounding-box minimum (voxelToWorld works on those indices)
                const int ix = ((bx << 3) | lx) - box.min[0], iy = ((by << 3) | ly) - box.min[1], iz = ((bz << 3) | lz) - box.min[2];
                const float ox = __fmul_rn(-(float)box.min[0], vs), oy = __fmul_rn(-(float)box.min[1], vs),
                            oz = __fmul_rn(-(float)box.min[2], vs);
                const unsigned long long row = kTriTable[cube];
                const int n = (int)(row >> 60);
                for (int t = 0; t < n; t += 3) {
                    float p[3][3];
#pragma unroll
                    for (int q = 0; q < 3; ++q) {
                        const int e = (int)((row >> (4 * (t + q))) & 0xF);
                        int a, b;
                        mc_edge(e, a, b);
                        int ax, ay, az, bx2, by2, bz2;
                        mc_corner(a, ax, ay, az);
                        mc_corner(b, bx2, by2, bz2);
                        const float va[3] = {__fsub_rn(__fmul_rn((float)(ix + ax), vs), ox), __fsub_rn(__fmul_rn((float)(iy + ay), vs), oy),
                                             __fsub_rn(__fmul_rn((float)(iz + az), vs), oz)};
                        const float vb[3] = {__fsub_rn(__fmul_rn((float)(ix + bx2), vs), ox), __fsub_rn(__fmul_rn((float)(iy + by2), vs), oy),
                                             __fsub_rn(__fmul_rn((float)(iz + bz2), vs), oz)};
                        float da = 0.f, db = 0.f;
#pragma unroll
                        for (int c = 0; c < 8; ++c) { da = c == a ? d[c] : da; db = c == b ? d[c] : db; }
                        mc_interpolate(da, db, va, vb, p[q]);
                    }
                    const bool ne01 = p[0][0] != p[1][0] || p[0][1] != p[1][1] || p[0][2] != p[1][2];
                    const bool ne02 = p[0][0] != p[2][0] || p[0][1] != p[2][1] || p[0][2] != p[2][2];
                    const bool ne12 = p[1][0] != p[2][0] || p[1][1] != p[2][1] || p[1][2] != p[2][2];
                    if (ne01 && ne02 && ne12) {
                        if (EMIT) {
#pragma unroll
                            for (int k = 0; k < 5; ++k)
                                if (k == ntri) {
#pragma unroll
                                    for (int q = 0; q < 9; ++q) tri[k][q] = p[q / 3][q % 3];
                                    tpos[k] = t / 3;
                                }
                        }
                        ++ntri;
                    }
                }
            }
            // warp-aggregated cursor
            int incl = ntri;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, incl, off);
                if (lane >= off) incl += o;
            }
            const int wtot = __shfl_sync(0xffffffffu, incl, 31);
            if (wtot == 0) continue;
            unsigned long long base = 0;
            if (lane == 31) base = atomicAdd(total, (unsigned long long)wtot);
            base = __shfl_sync(0xffffffffu, base, 31);
            if (EMIT && ntri > 0) {
                const long long lin = ((long long)(((bz << 3) | lz) - box.min[2]) * box.dim[1] + (((by << 3) | ly) - box.min[1])) * box.dim[0] +
                                      (((bx << 3) | lx) - box.min[0]);
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const long long o = (long long)base + (incl - ntri) + k;
                    if (k < ntri && o < max_tris) {
                        keys[o] = (unsigned long long)lin * 8ull + (unsigned long long)tpos[k];
#pragma unroll
                        for (int q = 0; q < 9; ++q) verts[9 * o + q] = tri[k][q];
                    }
                }
            }
        }
    }
}

}  // namespace gsdf
