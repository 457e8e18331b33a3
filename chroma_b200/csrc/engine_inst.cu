// engine_inst.cu — the two-level ("instanced") geometry: device-side detector assembly, the
// kernels that search it (first-hit intersection, fused persistent propagate) and the lazily
// materialised flat arrays of the GPUGeometry API.  Layout and search: inst.cuh.
//
// Replaces, for detectors made of placed solids, chroma/geometry.py:360-367 (host flattening of every
// placement, chroma/detector.py:37-60 add_solid / add_pmt) + chroma/gpu/geometry.py:191-255 (upload of
// the flattened mesh and its BVH): the host hands over each distinct solid once and one
// (rotation, displacement) per placement; boxes, trees and placement records are made here.
// Search variants of this translation unit (r2f A/B on one B200, ms per 1e7-photon C3 bomb / per
// 2.5e6-photon C4 event; (two-pass group step, leaf spread) = 11 / 00 / 01 / 10):
//   flat C3 26.37 / 27.29 / 26.65 / 27.56    two-level C3 25.38 / 24.36 / 25.22 / 24.76
//   flat C4 19.94 / 18.33 / 19.21 / 18.86    two-level C4 23.40 / 22.05 / 24.29 / 22.63
// -> 11 only where the search of a huge flat tree dominates (engine_wide.cu); 00 here.
// r2j, with the predicated pushes of the two-pass step (WIDE_FAST_PUSH): two-level C3 24.31 -> 23.60 ms
// with the two-pass step here (10); the out-of-line-RNG builds stay 00 (C4 18.28 vs 18.47, 21.66 vs 21.87).
#ifndef WIDE_TWO_PASS
#define WIDE_TWO_PASS 1
#endif
#ifndef WIDE_LEAF_SPREAD
#define WIDE_LEAF_SPREAD 0
#endif
#include <math.h>
#include <stdlib.h>
#include <algorithm>
#include "physics.cuh"
#include "inst.cuh"
#include "inst_prop_kernel.cuh"

static uint32_t env_u32i(const char *name, uint32_t dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (uint32_t)strtoul(s, nullptr, 10);
}

// ---------------------------------------------------------------------------------------
// build kernels
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t ispread21(uint32_t v) {
    uint64_t x = v & 0x1FFFFFu;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

struct TemplateInfo {       // one per template, device copy
    float origin[3];        // template frame origin in the solid's own coordinates (box corner - margin)
    float scale;            // grid quantum
    uint32_t pad;           // quanta of padding around every box
    uint32_t voff, toff;    // first vertex / first triangle in the concatenated arrays
    uint32_t ntri;
};

// triangle records in the template frame, Morton codes and padded boxes on the template grid
__global__ void ik_template_kernel(uint64_t nt, const uint32_t *tri, const float *vert, const uint32_t *mcode,
                                   const uint32_t *tri_tmpl, const TemplateInfo *info, float4 *recs,
                                   uint64_t *codes, uint32_t *boxes) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < nt; t += stride) {
        const uint32_t tm = tri_tmpl[t];
        const TemplateInfo ti = info[tm];
        float v[3][3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float *p = vert + 3 * ((size_t)ti.voff + tri[3 * t + k]);
#pragma unroll
            for (int a = 0; a < 3; a++) v[k][a] = p[a] - ti.origin[a];
        }
        float4 *q = recs + 3 * t;
        q[0] = make_float4(v[0][0], v[0][1], v[0][2], v[1][0]);
        q[1] = make_float4(v[1][1], v[1][2], v[2][0], v[2][1]);
        q[2] = make_float4(v[2][2], __uint_as_float(mcode[t]), __uint_as_float((uint32_t)t - ti.toff),
                           __uint_as_float(tm));
        const float is = 1.0f / ti.scale;
        uint32_t box[3];
        uint64_t code = 0;
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float lo = fminf(v[0][a], fminf(v[1][a], v[2][a])), hi = fmaxf(v[0][a], fmaxf(v[1][a], v[2][a]));
            const float ql = floorf(lo * is) - (float)ti.pad, qh = ceilf(hi * is) + (float)ti.pad;
            box[a] = (uint32_t)fminf(fmaxf(ql, 0.0f), 65535.0f) | (uint32_t)fminf(fmaxf(qh, 0.0f), 65535.0f) << 16;
            const float c = (v[0][a] + v[1][a] + v[2][a]) * (1.0f / 3.0f) * is * 32.0f;
            code |= ispread21((uint32_t)fminf(fmaxf(c, 0.0f), 2097151.0f)) << a;
        }
        boxes[3 * t] = box[0]; boxes[3 * t + 1] = box[1]; boxes[3 * t + 2] = box[2];
        codes[t] = code;
    }
}

// records of one template into Morton order; triangle id inside the solid -> slot
__global__ void ik_permute_kernel(const float4 *recs, const uint32_t *perm, uint32_t off, uint32_t n, float4 *out,
                                  int32_t *tri_slot) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t s = perm[k];
    const float4 *p = recs + 3 * ((size_t)off + s);
    float4 *q = out + 3 * ((size_t)off + k);
    q[0] = p[0]; q[1] = p[1]; q[2] = p[2];
    tri_slot[off + s] = (int32_t)(off + k);
}

// append a tree to the common group array: child groups move by goff, leaf slots by soff
__global__ void ik_offset_groups_kernel(const uint4 *src, uint32_t nrec, uint32_t goff, uint32_t soff, uint4 *dst) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nrec) return;
    uint4 r = src[i];
    if (r.w != WIDE_EMPTY) {
        if (r.w & WIDE_LEAF) r.w = (r.w & 0xF0000000u) | ((r.w & 0x0FFFFFFFu) + soff);
        else r.w += goff;
    }
    dst[i] = r;
}

struct PlacementIn {     // host -> device, add order
    float r[9];          // rotation, row major: world = R v + d
    float d[3];          // displacement of the TEMPLATE FRAME origin: d + R * origin
    uint32_t tmpl;
};

// world box of every placement: all template vertices through its transform (one block per placement)
__global__ void ik_placement_box_kernel(const PlacementIn *pl, const TemplateInfo *info, const float *vert,
                                        const uint32_t *tmpl_nvert, float *lo_hi /* [n][6] */) {
    const PlacementIn P = pl[blockIdx.x];
    const TemplateInfo ti = info[P.tmpl];
    const uint32_t nv = tmpl_nvert[P.tmpl];
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = threadIdx.x; i < nv; i += blockDim.x) {
        const float *p = vert + 3 * ((size_t)ti.voff + i);
        const float x = p[0] - ti.origin[0], y = p[1] - ti.origin[1], z = p[2] - ti.origin[2];
#pragma unroll
        for (int a = 0; a < 3; a++) {
            const float w = P.r[3 * a] * x + P.r[3 * a + 1] * y + P.r[3 * a + 2] * z + P.d[a];
            lo[a] = fminf(lo[a], w); hi[a] = fmaxf(hi[a], w);
        }
    }
    __shared__ float s_lo[3][8], s_hi[3][8];
#pragma unroll
    for (int a = 0; a < 3; a++) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0) { s_lo[a][threadIdx.x >> 5] = lo[a]; s_hi[a][threadIdx.x >> 5] = hi[a]; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int a = threadIdx.x;
        float l = INFINITY, h = -INFINITY;
        for (unsigned w = 0; w < (blockDim.x >> 5); w++) { l = fminf(l, s_lo[a][w]); h = fmaxf(h, s_hi[a][w]); }
        lo_hi[6 * (size_t)blockIdx.x + a] = l;
        lo_hi[6 * (size_t)blockIdx.x + 3 + a] = h;
    }
}

// records of the top-tree leaves in leaf order: item k = (placement, root group to enter it through)
__global__ void ik_records_kernel(const PlacementIn *pl, const TemplateInfo *info, const uint32_t *perm,
                                  const uint32_t *item_place, const uint32_t *item_root, const uint32_t *tri_base,
                                  uint32_t nitems, float4 *rec) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nitems) return;
    const uint32_t it = perm[k], o = item_place[it];
    const PlacementIn P = pl[o];
    float4 *q = rec + 4 * (size_t)k;
    q[0] = make_float4(P.r[0], P.r[1], P.r[2], P.d[0]);
    q[1] = make_float4(P.r[3], P.r[4], P.r[5], P.d[1]);
    q[2] = make_float4(P.r[6], P.r[7], P.r[8], P.d[2]);
    q[3] = make_float4(info[P.tmpl].scale, __uint_as_float(item_root[it]), __uint_as_float(tri_base[o]),
                       __uint_as_float(o));
}
__global__ void ik_tmpl_off_kernel(const PlacementIn *pl, const TemplateInfo *info, uint32_t n, uint32_t *tmpl_off_of_orig) {
    uint32_t o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o < n) tmpl_off_of_orig[o] = info[pl[o].tmpl].toff;
}

// ---------------------------------------------------------------------------------------
// chroma_geometry_create_instanced
// ---------------------------------------------------------------------------------------
#define ICK(expr)                                                                          \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            chroma_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            rc = CHROMA_ERR_CUDA;                                                          \
            goto done;                                                                     \
        }                                                                                  \
    } while (0)
#define ITRY(x)                                                                            \
    do {                                                                                   \
        rc = (x);                                                                          \
        if (rc) goto done;                                                                 \
    } while (0)

template <typename T>
static cudaError_t dev_upload(T **dptr, const T *src, size_t count, cudaStream_t st) {
    cudaError_t e = cudaMalloc((void **)dptr, (count ? count : 1) * sizeof(T));
    if (e != cudaSuccess) return e;
    if (count) return cudaMemcpyAsync(*dptr, src, count * sizeof(T), cudaMemcpyHostToDevice, st);
    return cudaSuccess;
}

extern "C" int chroma_geometry_create_instanced(chroma_ctx *c, const chroma_geometry_desc *od,
                                                const chroma_instances_desc *d, chroma_geometry **out) {
    CTX_BIND(c);
    ARG_CHECK(od && d && out);
    ARG_CHECK(d->ntemplates > 0 && d->ninstances > 0 && d->ninstances < (1ull << 28));
    ARG_CHECK(d->vertices && d->triangles && d->material_codes && d->vertex_offset && d->triangle_offset);
    ARG_CHECK(d->instance_template && d->rotations && d->displacements);
    const uint32_t ntm = d->ntemplates;
    const uint64_t ninst = d->ninstances;
    const uint64_t nv = d->vertex_offset[ntm], ntt = d->triangle_offset[ntm];
    ARG_CHECK(nv > 0 && ntt > 0 && ntt < (1ull << 28));
    int rc = 0;
    cudaStream_t st = c->stream;

    // ---- host: template frames, placements, global triangle ids ----
    std::vector<TemplateInfo> info(ntm);
    std::vector<uint32_t> tmpl_nvert(ntm), tri_tmpl(ntt);
    std::vector<double> tmin(3 * (size_t)ntm), text(ntm);
    for (uint32_t t = 0; t < ntm; t++) {
        const uint64_t v0 = d->vertex_offset[t], v1 = d->vertex_offset[t + 1];
        const uint64_t t0 = d->triangle_offset[t], t1 = d->triangle_offset[t + 1];
        ARG_CHECK(v1 > v0 && t1 > t0);
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (uint64_t i = v0; i < v1; i++)
            for (int a = 0; a < 3; a++) {
                const double x = d->vertices[3 * i + a];
                if (!(x == x) || fabs(x) > 1e30) { chroma_set_error("template %u: vertex %llu is not finite", t, (unsigned long long)(i - v0)); return CHROMA_ERR_ARG; }
                lo[a] = std::min(lo[a], x); hi[a] = std::max(hi[a], x);
            }
        for (uint64_t i = t0; i < t1; i++) {
            tri_tmpl[i] = t;
            for (int k = 0; k < 3; k++)
                if (d->triangles[3 * i + k] >= v1 - v0) { chroma_set_error("template %u: triangle %llu refers to a vertex out of range", t, (unsigned long long)(i - t0)); return CHROMA_ERR_ARG; }
        }
        double ext = std::max(hi[0] - lo[0], std::max(hi[1] - lo[1], hi[2] - lo[2]));
        if (!(ext > 0)) ext = 1.0;
        text[t] = ext;
        for (int a = 0; a < 3; a++) tmin[3 * t + a] = lo[a];
        info[t].voff = (uint32_t)v0; info[t].toff = (uint32_t)t0; info[t].ntri = (uint32_t)(t1 - t0);
        tmpl_nvert[t] = (uint32_t)(v1 - v0);
    }
    // how far from the origin the detector reaches: the template-frame ray carries ~ulp(reach) of
    // rounding from the subtraction o - d', which the padding of the template boxes must cover
    double reach = 0;
    for (uint64_t i = 0; i < ninst; i++) {
        ARG_CHECK(d->instance_template[i] < ntm);
        double m = 0;
        for (int a = 0; a < 3; a++) m = std::max(m, fabs((double)d->displacements[3 * i + a]));
        reach = std::max(reach, m + 2.0 * text[d->instance_template[i]]);
        const float *R = d->rotations + 9 * i;
        for (int a = 0; a < 3; a++)
            for (int b = 0; b < 3; b++) {
                double dot = 0;
                for (int k = 0; k < 3; k++) dot += (double)R[3 * a + k] * R[3 * b + k];
                if (fabs(dot - (a == b ? 1.0 : 0.0)) > 1e-4) {
                    chroma_set_error("placement %llu: not a pure rotation (R R^T differs from 1 by %g); the two-level "
                                     "tree needs length-preserving placements, use the flat geometry instead",
                                     (unsigned long long)i, fabs(dot - (a == b ? 1.0 : 0.0)));
                    return CHROMA_ERR_ARG;
                }
            }
    }
    const double ray_err = 16.0 * reach * 1.1920929e-7;   // 16 ulp of the largest coordinate
    for (uint32_t t = 0; t < ntm; t++) {
        // grid: 65534 quanta across extent + 2 margins, margin = (pad + 1) quanta
        uint32_t pad = 2;
        double scale = 0;
        for (int it = 0; it < 8; it++) {
            scale = text[t] / (65534.0 - 2.0 * (pad + 1));
            uint32_t need = 2 + (uint32_t)ceil(ray_err / scale);
            if (need > 1024) need = 1024;
            if (need <= pad) break;
            pad = need;
        }
        info[t].scale = (float)scale;
        info[t].pad = pad;
        for (int a = 0; a < 3; a++) info[t].origin[a] = (float)(tmin[3 * t + a] - (pad + 1) * scale);
    }
    std::vector<PlacementIn> pl(ninst);
    std::vector<uint32_t> tri_base(ninst + 1);
    uint64_t acc = 0;
    for (uint64_t i = 0; i < ninst; i++) {
        const uint32_t t = d->instance_template[i];
        const float *R = d->rotations + 9 * i;
        for (int k = 0; k < 9; k++) pl[i].r[k] = R[k];
        for (int a = 0; a < 3; a++) {
            double x = d->displacements[3 * i + a];
            for (int k = 0; k < 3; k++) x += (double)R[3 * a + k] * (double)info[t].origin[k];
            pl[i].d[a] = (float)x;
        }
        pl[i].tmpl = t;
        tri_base[i] = (uint32_t)acc;
        acc += info[t].ntri;
    }
    if (acc >= (1ull << 31)) { chroma_set_error("%llu triangles after placement: triangle ids must fit 31 bits", (unsigned long long)acc); return CHROMA_ERR_ARG; }
    tri_base[ninst] = (uint32_t)acc;

    chroma_geometry *g = new chroma_geometry();
    g->ctx = c;
    g->instanced = true;
    g->tree_kind = 2;
    g->ntriangles = acc;
    g->nvertices = 0;
    g->nsolids = ninst;
    g->ntemplates = ntm;
    // device scratch
    float *d_vert = nullptr; uint32_t *d_tri = nullptr, *d_mcode = nullptr, *d_tri_tmpl = nullptr, *d_nvert = nullptr;
    TemplateInfo *d_info = nullptr; PlacementIn *d_pl = nullptr;
    float4 *d_recs = nullptr; uint64_t *d_codes = nullptr; uint32_t *d_boxes = nullptr;
    float *d_lohi = nullptr; uint64_t *d_tcodes = nullptr; uint32_t *d_tboxes = nullptr, *d_item_place = nullptr, *d_item_root = nullptr;
    std::vector<WideTreeOut> trees(ntm);
    WideTreeOut top;
    std::vector<float> lohi;
    std::vector<uint32_t> root_group(ntm);
    {
        ICK(dev_upload(&d_vert, d->vertices, nv * 3, st));
        ICK(dev_upload(&d_tri, d->triangles, ntt * 3, st));
        ICK(dev_upload(&d_mcode, d->material_codes, ntt, st));
        ICK(dev_upload(&d_tri_tmpl, tri_tmpl.data(), ntt, st));
        ICK(dev_upload(&d_nvert, tmpl_nvert.data(), (size_t)ntm, st));
        ICK(dev_upload(&d_info, info.data(), (size_t)ntm, st));
        ICK(dev_upload(&d_pl, pl.data(), (size_t)ninst, st));
        ICK(cudaMalloc(&d_recs, ntt * 48)); ICK(cudaMalloc(&d_codes, ntt * 8)); ICK(cudaMalloc(&d_boxes, ntt * 12));
        ik_template_kernel<<<grid_for(ntt, 256, c, 8), 256, 0, st>>>(ntt, d_tri, d_vert, d_mcode, d_tri_tmpl, d_info,
                                                                     d_recs, d_codes, d_boxes);
        c->launches++;
        ICK(cudaGetLastError());
        // ---- one tree per template, records into Morton order ----
        ICK(cudaMalloc(&g->d_tris, ntt * 48));
        ICK(cudaMalloc(&g->d_tri_slot, ntt * 4));
        const uint32_t leaf_max = std::min(8u, std::max(1u, env_u32i("CHROMA_WIDE_LEAF", 3)));
        for (uint32_t t = 0; t < ntm; t++) {
            ITRY(wide_tree_from_boxes(c, info[t].ntri, d_codes + info[t].toff, d_boxes + 3 * (size_t)info[t].toff,
                                      leaf_max, &trees[t]));
            ik_permute_kernel<<<(info[t].ntri + 255) / 256, 256, 0, st>>>(d_recs, trees[t].perm, info[t].toff,
                                                                            info[t].ntri, g->d_tris, g->d_tri_slot);
            c->launches++;
        }
        ICK(cudaGetLastError());
        // ---- template trees first in the common group array (their offsets are needed below) ----
        uint64_t total = 0;
        for (uint32_t t = 0; t < ntm; t++) { root_group[t] = (uint32_t)total; total += trees[t].ngroups; }
        // ---- placement boxes ----
        ICK(cudaMalloc(&d_lohi, ninst * 6 * sizeof(float)));
        ik_placement_box_kernel<<<(unsigned)ninst, 256, 0, st>>>(d_pl, d_info, d_vert, d_nvert, d_lohi);
        c->launches++;
        lohi.resize(ninst * 6);
        ICK(cudaMemcpyAsync(lohi.data(), d_lohi, ninst * 6 * sizeof(float), cudaMemcpyDeviceToHost, st));
        ICK(cudaStreamSynchronize(st));
        // ---- top-tree items: one per placement; a placement whose box dwarfs the typical one is entered
        // through the subtrees of its template tree `open_depth` levels down (see inst.cuh) ----
        std::vector<float> item_box;          // [nitems][6]
        std::vector<uint32_t> item_place, item_root;
        {
            std::vector<double> vol(ninst);
            for (uint64_t i = 0; i < ninst; i++)
                vol[i] = std::max(1e-9, (double)(lohi[6 * i + 3] - lohi[6 * i])) * std::max(1e-9, (double)(lohi[6 * i + 4] - lohi[6 * i + 1])) *
                         std::max(1e-9, (double)(lohi[6 * i + 5] - lohi[6 * i + 2]));
            std::vector<double> sorted(vol);
            std::nth_element(sorted.begin(), sorted.begin() + sorted.size() / 2, sorted.end());
            const double median = sorted[sorted.size() / 2];
            const int open_depth = (int)env_u32i("CHROMA_INST_OPEN", 3);
            std::vector<std::vector<uint4>> host_groups(ntm);
            auto template_groups = [&](uint32_t t) -> const std::vector<uint4> & {
                if (host_groups[t].empty()) {
                    host_groups[t].resize((size_t)trees[t].ngroups * 8);
                    cudaMemcpy(host_groups[t].data(), trees[t].groups, host_groups[t].size() * sizeof(uint4), cudaMemcpyDeviceToHost);
                }
                return host_groups[t];
            };
            for (uint64_t i = 0; i < ninst; i++) {
                const uint32_t t = pl[i].tmpl;
                const bool open = open_depth > 0 && ninst > 1 && vol[i] > 64.0 * median && trees[t].ngroups > 8;
                if (!open) {
                    item_place.push_back((uint32_t)i); item_root.push_back(root_group[t]);
                    item_box.insert(item_box.end(), lohi.begin() + 6 * i, lohi.begin() + 6 * i + 6);
                    continue;
                }
                const std::vector<uint4> &G = template_groups(t);
                // (group, depth left); a group is split further only if all its children are groups
                std::vector<std::pair<uint32_t, int>> todo{{0u, open_depth}};
                while (!todo.empty()) {
                    auto [grp, depth] = todo.back();
                    todo.pop_back();
                    const uint4 *rec = G.data() + (size_t)grp * 8;
                    bool all_inner = depth > 0;
                    for (int j = 0; j < 8 && all_inner; j++)
                        if (rec[j].w != WIDE_EMPTY && (rec[j].w & WIDE_LEAF)) all_inner = false;
                    if (all_inner) {
                        for (int j = 0; j < 8; j++)
                            if (rec[j].w != WIDE_EMPTY) todo.push_back({rec[j].w, depth - 1});
                        continue;
                    }
                    // world box of this subtree: the union of its records' boxes (template grid), its 8
                    // corners through the placement
                    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
                    for (int j = 0; j < 8; j++) {
                        if (rec[j].w == WIDE_EMPTY) continue;
                        const uint32_t w3[3] = {rec[j].x, rec[j].y, rec[j].z};
                        for (int a = 0; a < 3; a++) {
                            lo[a] = std::min(lo[a], (double)(w3[a] & 0xFFFFu) * info[t].scale);
                            hi[a] = std::max(hi[a], (double)(w3[a] >> 16) * info[t].scale);
                        }
                    }
                    float wb[6] = {INFINITY, INFINITY, INFINITY, -INFINITY, -INFINITY, -INFINITY};
                    for (int cnr = 0; cnr < 8; cnr++) {
                        const double v[3] = {(cnr & 1) ? hi[0] : lo[0], (cnr & 2) ? hi[1] : lo[1], (cnr & 4) ? hi[2] : lo[2]};
                        for (int a = 0; a < 3; a++) {
                            const double x = pl[i].r[3 * a] * v[0] + pl[i].r[3 * a + 1] * v[1] + pl[i].r[3 * a + 2] * v[2] + pl[i].d[a];
                            wb[a] = std::min(wb[a], (float)x); wb[3 + a] = std::max(wb[3 + a], (float)x);
                        }
                    }
                    // never outside the placement's own box
                    for (int a = 0; a < 3; a++) { wb[a] = std::max(wb[a], lohi[6 * i + a]); wb[3 + a] = std::min(wb[3 + a], lohi[6 * i + 3 + a]); }
                    item_place.push_back((uint32_t)i); item_root.push_back(root_group[t] + grp);
                    item_box.insert(item_box.end(), wb, wb + 6);
                }
            }
        }
        const uint64_t nitems = item_place.size();
        double wlo[3] = {1e300, 1e300, 1e300}, whi[3] = {-1e300, -1e300, -1e300};
        for (uint64_t i = 0; i < ninst; i++)
            for (int a = 0; a < 3; a++) {
                wlo[a] = std::min(wlo[a], (double)lohi[6 * i + a]);
                whi[a] = std::max(whi[a], (double)lohi[6 * i + 3 + a]);
            }
        double wext = std::max(whi[0] - wlo[0], std::max(whi[1] - wlo[1], whi[2] - wlo[2]));
        if (!(wext > 0)) wext = 1.0;
        const double tscale = wext / (65534.0 - 8.0);
        const float top_scale = (float)tscale;
        const float top_origin[3] = {(float)(wlo[0] - 4 * tscale), (float)(wlo[1] - 4 * tscale), (float)(wlo[2] - 4 * tscale)};
        std::vector<uint64_t> tcodes(nitems);
        std::vector<uint32_t> tboxes(3 * nitems);
        auto spread = [](uint32_t v) {
            uint64_t x = v & 0x1FFFFFu;
            x = (x | x << 32) & 0x1F00000000FFFFull; x = (x | x << 16) & 0x1F0000FF0000FFull;
            x = (x | x << 8) & 0x100F00F00F00F00Full; x = (x | x << 4) & 0x10C30C30C30C30C3ull;
            x = (x | x << 2) & 0x1249249249249249ull;
            return x;
        };
        for (uint64_t i = 0; i < nitems; i++) {
            uint64_t code = 0;
            for (int a = 0; a < 3; a++) {
                // two quanta of padding: the folded slab offset (wide.cuh) + the float box itself
                double ql = floor(((double)item_box[6 * i + a] - top_origin[a]) / top_scale) - 2.0;
                double qh = ceil(((double)item_box[6 * i + 3 + a] - top_origin[a]) / top_scale) + 2.0;
                ql = std::min(std::max(ql, 0.0), 65535.0); qh = std::min(std::max(qh, 0.0), 65535.0);
                tboxes[3 * i + a] = (uint32_t)ql | (uint32_t)qh << 16;
                double cc = (0.5 * ((double)item_box[6 * i + a] + item_box[6 * i + 3 + a]) - top_origin[a]) / top_scale * 32.0;
                code |= spread((uint32_t)std::min(std::max(cc, 0.0), 2097151.0)) << a;
            }
            tcodes[i] = code;
        }
        ICK(dev_upload(&d_tcodes, tcodes.data(), (size_t)nitems, st));
        ICK(dev_upload(&d_tboxes, tboxes.data(), 3 * (size_t)nitems, st));
        ITRY(wide_tree_from_boxes(c, nitems, d_tcodes, d_tboxes, 1, &top));
        // ---- one group array: the templates, then the top tree ----
        const uint32_t top_root = (uint32_t)total;
        total += top.ngroups;
        if (total >= (1ull << 28)) { chroma_set_error("too many groups"); rc = CHROMA_ERR_OVERFLOW; goto done; }
        ICK(cudaMalloc(&g->d_wgroups, total * 128));
        for (uint32_t t = 0; t < ntm; t++) {
            ik_offset_groups_kernel<<<(trees[t].ngroups * 8 + 255) / 256, 256, 0, st>>>(
                trees[t].groups, trees[t].ngroups * 8, root_group[t], info[t].toff, g->d_wgroups + (size_t)root_group[t] * 8);
            c->launches++;
        }
        ik_offset_groups_kernel<<<(top.ngroups * 8 + 255) / 256, 256, 0, st>>>(top.groups, top.ngroups * 8, top_root, 0u,
                                                                              g->d_wgroups + (size_t)top_root * 8);
        c->launches++;
        ICK(cudaGetLastError());
        // ---- leaf records in leaf order + lookup arrays ----
        ICK(dev_upload(&d_item_place, item_place.data(), (size_t)nitems, st));
        ICK(dev_upload(&d_item_root, item_root.data(), (size_t)nitems, st));
        ICK(dev_upload(&g->d_tri_base, tri_base.data(), (size_t)ninst + 1, st));
        ICK(cudaMalloc(&g->d_inst, nitems * 64));
        ICK(cudaMalloc(&g->d_tmpl_off_of_orig, ninst * 4));
        ik_records_kernel<<<(unsigned)((nitems + 255) / 256), 256, 0, st>>>(d_pl, d_info, top.perm, d_item_place, d_item_root,
                                                                            g->d_tri_base, (uint32_t)nitems, g->d_inst);
        ik_tmpl_off_kernel<<<(unsigned)((ninst + 255) / 256), 256, 0, st>>>(d_pl, d_info, (uint32_t)ninst, g->d_tmpl_off_of_orig);
        c->launches += 2;
        if (d->colors) ICK(dev_upload(&g->d_tmpl_colors, d->colors, ntt, st));
        {   // first flat vertex of every solid (no vertex de-duplication across solids)
            std::vector<uint32_t> vbase(ninst + 1);
            uint64_t va = 0;
            for (uint64_t i = 0; i < ninst; i++) { vbase[i] = (uint32_t)va; va += tmpl_nvert[d->instance_template[i]]; }
            if (va >= (1ull << 32)) va = 0;   // too many for 32-bit indices: flat arrays unavailable
            vbase[ninst] = (uint32_t)va;
            g->flat_nvertices = va;
            ICK(dev_upload(&g->d_vert_base, vbase.data(), (size_t)ninst + 1, st));
        }
        g->d_tmpl_vert = d_vert; d_vert = nullptr;
        g->d_tmpl_tri = d_tri; d_tri = nullptr;
        g->d_tmpl_mcode = d_mcode; d_mcode = nullptr;
        g->d_tmpl_nvert = d_nvert; d_nvert = nullptr;
        g->d_tmpl_info = d_info; d_info = nullptr;
        g->d_placements = d_pl; d_pl = nullptr;
        ICK(cudaStreamSynchronize(st));
        ICK(cudaGetLastError());
        g->nslots = ntt;
        g->inst_top_groups = top.ngroups;
        g->inst_top_leaves = (uint32_t)nitems;
        g->wide_levels = top.levels;
        g->wide_leaves = top.leaves;
        g->bytes += total * 128 + ntt * 52 + nitems * 64 + ninst * 12 + 4;
        InstView &V = g->iv;
        V.groups = g->d_wgroups; V.tris = g->d_tris; V.inst = g->d_inst; V.tri_base = g->d_tri_base;
        V.tmpl_off_of_orig = g->d_tmpl_off_of_orig; V.tri_slot = g->d_tri_slot; V.top_root = top_root;
        V.ngroups = (uint32_t)total; V.ninst = (uint32_t)ninst; V.ntriangles = (uint32_t)acc; V.nslots = (uint32_t)ntt;
        V.top_origin = make_float3(top_origin[0], top_origin[1], top_origin[2]);
        V.top_scale = top_scale;
        {   // wide.cuh wide_origin_unsearchable: the finest grid (a template's) sets the bound
            float smin = top_scale;
            for (size_t t = 0; t < info.size(); t++) smin = std::min(smin, info[t].scale);
            V.guard_lim = WIDE_GUARD_QUANTA * smin;
        }
        // the same arrays through the WideView the group step reads (rank unused: no tie break here)
        g->wide = WideView{};
        g->wide.groups = g->d_wgroups; g->wide.tris = g->d_tris; g->wide.rank = nullptr;
        g->wide.ngroups = (uint32_t)total; g->wide.nslots = (uint32_t)ntt;
        g->wide.world_origin = V.top_origin; g->wide.world_scale = top_scale;
        ITRY(chroma_build_optics(c, g, od));
        g->view.tris = g->d_tris;           // template records (material codes for callers that only need those)
        g->view.tri_slot = nullptr;
        g->view.ntriangles = (uint32_t)acc;
        g->view.nslots = (uint32_t)ntt;
        g->view.world_origin = V.top_origin; g->view.world_scale = top_scale;
        g->view.inst_tri_base = g->d_tri_base;
        g->view.ninst = (uint32_t)ninst;
        g->wview = g->view;
    }
done:
    {
        void *tmp[] = {d_vert, d_tri, d_mcode, d_tri_tmpl, d_nvert, d_info, d_pl, d_recs, d_codes, d_boxes, d_lohi,
                       d_tcodes, d_tboxes, d_item_place, d_item_root, top.groups, top.perm};
        for (void *p : tmp) if (p) cudaFree(p);
        for (auto &t : trees) { if (t.groups) cudaFree(t.groups); if (t.perm) cudaFree(t.perm); }
    }
    if (rc) {
        chroma_geometry_destroy(g);
        return rc;
    }
    *out = g;
    return 0;
}

// ---------------------------------------------------------------------------------------
// first-hit intersection on the two-level tree (mesh.h:123-151 semantics + the triangle index)
// ---------------------------------------------------------------------------------------
#define IISECT_THREADS 128

// global triangle id of a hit: first triangle of the placement + triangle id inside the solid
__device__ __forceinline__ int inst_triangle_id(const InstView &v, int inst, int slot) {
    const uint32_t base = __float_as_uint(__ldg(v.inst + (size_t)inst * 4 + 3).z);
    return (int)(base + __float_as_uint(__ldg(v.tris + (size_t)slot * 3 + 2).z));
}

template <bool COUNT>
__global__ void __launch_bounds__(IISECT_THREADS)
inst_intersect_kernel(InstView v, WideView w, uint64_t n, const float *__restrict__ pos,
                      const float *__restrict__ dir, const int32_t *__restrict__ last_hit,
                      int32_t *__restrict__ out_tri, float *__restrict__ out_dist,
                      uint32_t *__restrict__ out_counts, uint32_t *status, unsigned long long *head, int refill,
                      int iters, int leaf_thresh) {
    extern __shared__ float4 smem_raw[];
    uint2 *s_stack = (uint2 *)smem_raw;
    float *s_side = (float *)(s_stack + WIDE_SSTACK * IISECT_THREADS);
    uint2 l_stack[WIDE_STACK - WIDE_SSTACK];
    const WStackRef stk = wstk_make(s_stack + threadIdx.x, l_stack, IISECT_THREADS);
    ISide sd;
    sd.p = s_side + threadIdx.x; sd.nthreads = IISECT_THREADS;
    WSpread spr;
    {
        unsigned char *sb = (unsigned char *)(s_side + ISIDE_WORDS * IISECT_THREADS) + (threadIdx.x >> 5) * WSPREAD_BYTES_PER_WARP;
        spr.res = (float *)sb; spr.own = sb + WSPREAD_PAIRS * 4;
    }
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    bool have = false, exhausted = false;
    unsigned overflow = 0u;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1);
    uint64_t id = 0;
    uint32_t ng = 0, nt = 0;
    ITrav tv;
    tv.cur = WIDE_NONE; tv.sp = -1; tv.hit = -1; tv.hit_inst = -1; tv.inst = -1; tv.dist = 0.f;
    while (true) {
        const bool work = have && !itrav_done(tv);
        const unsigned mW = __ballot_sync(0xffffffffu, work);
        const unsigned mH = __ballot_sync(0xffffffffu, have);
        if (mH == 0u && exhausted) break;
        const int nidle = __popc(mH & ~mW) + (exhausted ? 0 : 32 - __popc(mH));
        if (mW == 0u || nidle >= refill) {
            if (have && !work) {
                out_tri[id] = tv.hit >= 0 ? inst_triangle_id(v, tv.hit_inst, tv.hit) : -1;
                out_dist[id] = tv.hit >= 0 ? tv.dist : -1.0f;
                if (COUNT) { out_counts[2 * id] = ng; out_counts[2 * id + 1] = nt; }
                have = false;
            }
            if (!exhausted) {
                const unsigned hv = __ballot_sync(0xffffffffu, have);
                const int nfree = 32 - __popc(hv);
                if (nfree > 0) {
                    unsigned long long base = 0;
                    if (lane == 0) base = atomicAdd(head, (unsigned long long)nfree);
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (base + (unsigned long long)nfree >= n) exhausted = true;
                    if (!have) {
                        unsigned long long i = base + (unsigned long long)__popc(~hv & lt_mask);
                        if (i < n) {
                            id = i;
                            o = f3(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
                            d = f3(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
                            float nrm = sqrtf(ref_dot(d, d));   // mesh.h:141
                            d = f3(__fdividef(d.x, nrm), __fdividef(d.y, nrm), __fdividef(d.z, nrm));
                            int li = -1, ls = -1;
                            if (last_hit) inst_locate(v, last_hit[id], li, ls);
                            sd.i(6) = ls; sd.i(7) = li;
                            itrav_begin(tv, v, o, d);
                            if (wide_origin_unsearchable(o, v.guard_lim)) tv.cur = WIDE_NONE;   // no hit (wide.cuh)
                            ng = nt = 0;
                            have = true;
                        }
                    }
                }
            }
            continue;
        }
        for (int it = 0; it < iters; it++)
            itrav_round<COUNT>(tv, have, v, w, o, d, stk, sd, spr, leaf_thresh, overflow, ng, nt);
    }
    if (overflow) atomicOr(status, 1u);
}

int chroma_inst_intersect(chroma_geometry *g, uint64_t n, const float *pos, const float *dir,
                          const int32_t *last_hit, int32_t *out_tri, float *out_dist, uint32_t *out_counts) {
    chroma_ctx *c = g->ctx;
    if (n == 0) return 0;
    uint32_t *status = c->d_scratch + 2;
    CU_CHECK(cudaMemsetAsync(status, 0, 4, c->stream));
    unsigned long long *head = (unsigned long long *)(c->d_scratch + 10);
    CU_CHECK(cudaMemsetAsync(head, 0, 8, c->stream));
    const size_t smem = (size_t)IISECT_THREADS * (WIDE_SSTACK * sizeof(uint2) + ISIDE_WORDS * sizeof(float)) +
                        (INST_LEAF_SPREAD ? (IISECT_THREADS / 32) * WSPREAD_BYTES_PER_WARP : 0);
    const int refill = std::max(1, c->knobs.isect_refill), iters = std::max(1, c->knobs.isect_iters),
              lt = std::max(1, c->knobs.leaf_thresh);
    int occ = 0;
    if (out_counts) {
        CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, inst_intersect_kernel<true>, IISECT_THREADS, smem));
        inst_intersect_kernel<true><<<grid_for(n, IISECT_THREADS, c, (unsigned)std::max(occ, 1)), IISECT_THREADS, smem, c->stream>>>(
            g->iv, g->wide, n, pos, dir, last_hit, out_tri, out_dist, out_counts, status, head, refill, iters, lt);
    } else {
        CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, inst_intersect_kernel<false>, IISECT_THREADS, smem));
        inst_intersect_kernel<false><<<grid_for(n, IISECT_THREADS, c, (unsigned)std::max(occ, 1)), IISECT_THREADS, smem, c->stream>>>(
            g->iv, g->wide, n, pos, dir, last_hit, out_tri, out_dist, nullptr, status, head, refill, iters, lt);
    }
    KERNEL_CHECK(c);
    return 0;
}

// ---------------------------------------------------------------------------------------
// fused persistent propagate kernel on the two-level tree: engine_wide.cu wide_propagate_kernel with
// the search of inst.cuh.  After a hit the triangle's normal is rotated into the world frame
// (world = R v + d' => n_world = R n_template); everything else (draw order, physics, refill from
// the queue head, staged upload) is the flat kernel's.
// ---------------------------------------------------------------------------------------
int chroma_inst_prop_occupancy(chroma_ctx *c, bool ool, size_t *smem, int *occ) {
    *smem = (size_t)IPROP_THREADS * (WIDE_SSTACK * sizeof(uint2) + ISIDE_WORDS * sizeof(float)) +
            (INST_LEAF_SPREAD ? (IPROP_THREADS / 32) * WSPREAD_BYTES_PER_WARP : 0);
    if (ool) return chroma_inst_ool_occupancy(c, *smem, occ);
    if (!c->inst_attr_set) {
        CU_CHECK(cudaFuncSetAttribute(inst_propagate_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        CU_CHECK(cudaFuncSetAttribute(inst_propagate_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        c->inst_attr_set = true;
    }
    CU_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(occ, inst_propagate_kernel<false>, IPROP_THREADS, *smem));
    if (*occ < 1) *occ = 1;
    return 0;
}

void chroma_inst_prop_launch(chroma_geometry *g, chroma_photons *p, const WPropParams &prm, bool ool, bool staged,
                             unsigned grid, size_t smem) {
    chroma_ctx *c = g->ctx;
    if (ool) { chroma_inst_ool_launch(g, p, prm, staged, grid, smem); return; }
    if (staged) inst_propagate_kernel<true><<<grid, IPROP_THREADS, smem, c->stream>>>(g->wview, g->iv, g->wide, p->v, prm);
    else inst_propagate_kernel<false><<<grid, IPROP_THREADS, smem, c->stream>>>(g->wview, g->iv, g->wide, p->v, prm);
}

// ---------------------------------------------------------------------------------------
// flat arrays of the GPUGeometry API (.vertices .triangles .material_codes .colors .solid_id_map),
// assembled on the device the first time they are asked for: what chroma/geometry.py:360-367 computes
// on the host (world vertex = R v + d in float32 here, float64 rounded to float32 there: equal to an ulp)
// ---------------------------------------------------------------------------------------
__global__ void ik_flat_vertices_kernel(const PlacementIn *pl, const TemplateInfo *info, const float *vert,
                                        const uint32_t *tmpl_nvert, const uint32_t *vbase, float *out) {
    const PlacementIn P = pl[blockIdx.x];
    const TemplateInfo ti = info[P.tmpl];
    const uint32_t nv = tmpl_nvert[P.tmpl], base = vbase[blockIdx.x];
    for (uint32_t i = threadIdx.x; i < nv; i += blockDim.x) {
        const float *p = vert + 3 * ((size_t)ti.voff + i);
        const float x = p[0] - ti.origin[0], y = p[1] - ti.origin[1], z = p[2] - ti.origin[2];
#pragma unroll
        for (int a = 0; a < 3; a++)
            out[3 * ((size_t)base + i) + a] = P.r[3 * a] * x + P.r[3 * a + 1] * y + P.r[3 * a + 2] * z + P.d[a];
    }
}
// which: 2 triangles (3 words), 3 material codes, 4 colors, 5 solid id
__global__ void ik_flat_triangles_kernel(const PlacementIn *pl, const TemplateInfo *info, const uint32_t *tri,
                                         const uint32_t *per_tri, const uint32_t *vbase, const uint32_t *tri_base,
                                         int which, uint32_t *out) {
    const PlacementIn P = pl[blockIdx.x];
    const TemplateInfo ti = info[P.tmpl];
    const uint32_t base = tri_base[blockIdx.x], vb = vbase[blockIdx.x];
    for (uint32_t i = threadIdx.x; i < ti.ntri; i += blockDim.x) {
        const size_t s = (size_t)ti.toff + i, o = (size_t)base + i;
        if (which == 2) {
            out[3 * o] = tri[3 * s] + vb; out[3 * o + 1] = tri[3 * s + 1] + vb; out[3 * o + 2] = tri[3 * s + 2] + vb;
        } else if (which == 5) {
            out[o] = blockIdx.x;
        } else {
            out[o] = per_tri ? per_tri[s] : 0u;
        }
    }
}

int chroma_inst_flat_array(chroma_geometry *g, int which, void **dptr, uint64_t *count) {
    chroma_ctx *c = g->ctx;
    if (which == 0) { *dptr = nullptr; *count = 0; return 0; }   // no reference-format node array
    if (which < 1 || which > 5) { chroma_set_error("unknown geometry array %d", which); return CHROMA_ERR_ARG; }
    const uint64_t n = which == 1 ? g->flat_nvertices : g->ntriangles;
    if (which <= 2 && g->flat_nvertices == 0) {
        chroma_set_error("the flattened mesh has more than 2^32 vertices");
        return CHROMA_ERR_OVERFLOW;
    }
    if (!g->d_flat[which]) {
        const size_t words = (which <= 2 ? 3 : 1) * n;
        CU_CHECK(cudaMalloc(&g->d_flat[which], words * 4));
        g->bytes += words * 4;
        const PlacementIn *pl = (const PlacementIn *)g->d_placements;
        const TemplateInfo *info = (const TemplateInfo *)g->d_tmpl_info;
        if (which == 1)
            ik_flat_vertices_kernel<<<(unsigned)g->nsolids, 256, 0, c->stream>>>(pl, info, g->d_tmpl_vert, g->d_tmpl_nvert,
                                                                                  g->d_vert_base, (float *)g->d_flat[1]);
        else
            ik_flat_triangles_kernel<<<(unsigned)g->nsolids, 256, 0, c->stream>>>(
                pl, info, g->d_tmpl_tri, which == 3 ? g->d_tmpl_mcode : (which == 4 ? g->d_tmpl_colors : nullptr),
                g->d_vert_base, g->d_tri_base, which, (uint32_t *)g->d_flat[which]);
        KERNEL_CHECK(c);
    }
    *dptr = g->d_flat[which];
    *count = n;
    return 0;
}
