// engine_render.cu — headless alpha-blended ray cast of the geometry.
//
// Replaces chroma/cuda/render.cu:11-181 (`render`, `get_color`), chroma/cuda/transform.cu:10-49
// (`translate`, `rotate`, `rotate_around_point`) and their host chroma/gpu/render.py (GPURays).
//
// What the reference computes per ray: every triangle the ray hits is offered, in the tree's
// visiting order, to a list of the `alpha_depth` nearest hits that is kept sorted with a
// lower-bound search (sorting.h:62-98): a hit goes in FRONT of entries at the same distance, and a
// full list drops its last entry.  The list it ends with is therefore
//      the first alpha_depth of all hits ordered by (distance, later visit first),
// with entries carried over from a previous render (keep_last_render) behind new hits at the same
// distance — an order that does not depend on how the hits were found.  The reference walks every
// box the ray touches, however far; here the wide tree (wide.cuh) is searched near-first with the
// list's last distance as the pruning bound once the list is full, and ties are resolved by the
// per-slot reference visiting rank that the nearest-hit search already uses.  The blend and the
// colour arithmetic are the reference's expressions, compiled with its options (--use_fast_math).
#include "physics.cuh"
#include "wide.cuh"

struct chroma_rays {
    chroma_ctx *ctx = nullptr;
    uint64_t n = 0;
    uint32_t max_alpha_depth = 0;
    float *d_pos = nullptr, *d_dir = nullptr;   // float3[n]
    float *d_dx = nullptr;                       // [n][max_alpha_depth] hit distances
    float4 *d_color = nullptr;                   // [n][max_alpha_depth] (r, g, b)*cos, alpha
    uint32_t *d_dxlen = nullptr;                 // [n] entries in use
};

#define RENDER_BLOCK 128
#define RENDER_OLD 0x80000000u   // list source: entry i of the previous render

// render.cu:11-33
__device__ __forceinline__ float4 render_color(const float3 &direction, const float3 &v0, const float3 &v1,
                                               const float3 &v2, unsigned rgba) {
    float3 v01 = f3(v1.x - v0.x, v1.y - v0.y, v1.z - v0.z);
    float3 v12 = f3(v2.x - v1.x, v2.y - v1.y, v2.z - v1.z);
    float3 c = f3(v01.y * v12.z - v01.z * v12.y, v01.z * v12.x - v01.x * v12.z, v01.x * v12.y - v01.y * v12.x);
    float len = sqrtf(c.x * c.x + c.y * c.y + c.z * c.z);
    float3 sn = f3(c.x / len, c.y / len, c.z / len);
    float cos_theta = sn.x * (-direction.x) + sn.y * (-direction.y) + sn.z * (-direction.z);
    if (cos_theta < 0.0f) cos_theta = -cos_theta;
    unsigned a0 = 0xff & (rgba >> 24), r0 = 0xff & (rgba >> 16), g0 = 0xff & (rgba >> 8), b0 = 0xff & rgba;
    float alpha = (255 - a0) / 255.0f;
    return make_float4(r0 * cos_theta, g0 * cos_theta, b0 * cos_theta, alpha);
}

// does list entry (dist e, source es) stay in front of a new hit (dist d, rank of slot s)?
__device__ __forceinline__ bool render_keeps_front(float e, unsigned es, float d, unsigned srank, const uint32_t *rank) {
    if (e < d) return true;
    if (e > d) return false;
    if (es & RENDER_OLD) return false;          // new hits go in front of carried-over entries
    return __ldg(rank + es) > srank;            // among this render's hits the later visit is in front
}

__global__ void __launch_bounds__(RENDER_BLOCK)
render_kernel(uint64_t nrays, const float *pos, const float *dir, WideView w, const uint32_t *colors,
              unsigned alpha_depth, uint32_t *pixels, float *dx_all, uint32_t *dxlen, float4 *color_all,
              unsigned bg_color, uint32_t *status) {
    extern __shared__ uint2 smem[];
    uint2 *sstack = smem;                                     // [WIDE_SSTACK][RENDER_BLOCK]
    float *ld = (float *)(smem + WIDE_SSTACK * RENDER_BLOCK); // [alpha_depth][RENDER_BLOCK]
    unsigned *ls = (unsigned *)(ld + alpha_depth * RENDER_BLOCK);
    uint2 spill[WIDE_STACK - WIDE_SSTACK];
    const uint64_t id = (uint64_t)blockIdx.x * RENDER_BLOCK + threadIdx.x;
    if (id >= nrays) return;
    ld += threadIdx.x;
    ls += threadIdx.x;
    const WStackRef k = wstk_make(sstack + threadIdx.x, spill, RENDER_BLOCK);

    const float3 o = f3(pos[3 * id], pos[3 * id + 1], pos[3 * id + 2]);
    const float3 d = f3(dir[3 * id], dir[3 * id + 1], dir[3 * id + 2]);
    float *dx = dx_all + id * alpha_depth;
    float4 *color_a = color_all + id * alpha_depth;
    unsigned n = dxlen[id];
    if (n > alpha_depth) n = alpha_depth;
    for (unsigned i = 0; i < n; i++) {
        ld[i * RENDER_BLOCK] = dx[i];
        ls[i * RENDER_BLOCK] = RENDER_OLD | i;
    }

    WTrav t;
    wtrav_begin(t, w, o, d, -1);
    if (n == alpha_depth) t.dist = ld[(alpha_depth - 1) * RENDER_BLOCK];
    RayCtx rc;
    rc.o = o;
    rc.d = d;
    unsigned overflow = 0;
    uint32_t unused = 0;
    while (t.cur != WIDE_NONE) {
        if (t.cur & WIDE_LEAF) {
            const int first = (int)(t.cur & 0x0FFFFFFFu);
            const int cnt = (int)((t.cur >> 28) & 7u) + 1;
            for (int j = 0; j < cnt; j++) {
                const int slot = first + j;
                const float4 *p = w.tris + (size_t)slot * 3;
                const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
                float tt;
                if (!triangle_hit(rc, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), f3(b.z, b.w, c.x), tt)) continue;
                // position = entries that stay in front; a full list only takes what beats its last entry
                const unsigned srank = __ldg(w.rank + slot);
                int j2 = (int)n;
                while (j2 > 0 && !render_keeps_front(ld[(j2 - 1) * RENDER_BLOCK], ls[(j2 - 1) * RENDER_BLOCK], tt, srank, w.rank)) j2--;
                if ((unsigned)j2 <= alpha_depth - 1) {
                    const int last = (int)min(n, alpha_depth - 1);
                    for (int m = last; m > j2; m--) {
                        ld[m * RENDER_BLOCK] = ld[(m - 1) * RENDER_BLOCK];
                        ls[m * RENDER_BLOCK] = ls[(m - 1) * RENDER_BLOCK];
                    }
                    ld[j2 * RENDER_BLOCK] = tt;
                    ls[j2 * RENDER_BLOCK] = (unsigned)slot;
                }
                if (n < alpha_depth) n++;
                if (n == alpha_depth) t.dist = ld[(alpha_depth - 1) * RENDER_BLOCK];
            }
            wtrav_pop(t, k);
        } else {
            wtrav_group_step<false>(t, w, k, overflow, unused);
        }
    }
    if (overflow) atomicOr(status, 1u);

    if (n < 1) {                       // render.cu:61-64,136-139: nothing hit, list untouched
        pixels[id] = bg_color;
        return;
    }
    // the lists go back to global memory in the reference's layout.  Entries only ever move towards
    // the back, so writing from the back never overwrites a carried-over colour that is still needed
    for (int i = (int)n - 1; i >= 0; i--) {
        const unsigned src = ls[i * RENDER_BLOCK];
        float4 col;
        if (src & RENDER_OLD) {
            col = color_a[src & ~RENDER_OLD];
        } else {
            const float4 *p = w.tris + (size_t)src * 3;
            const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
            col = render_color(d, f3(a.x, a.y, a.z), f3(a.w, b.x, b.y), f3(b.z, b.w, c.x),
                               __ldg(colors + __float_as_uint(c.z)));
        }
        color_a[i] = col;
        dx[i] = ld[i * RENDER_BLOCK];
    }
    dxlen[id] = n;

    // render.cu:143-178
    float scale = 1.0f;
    float fr = 0.0;
    float fg = 0.0;
    float fb = 0.0;
    for (int i = 0; i < (int)n; i++) {
        const float4 ci = color_a[i];
        float alpha = ci.w;
        fr += scale * ci.x * alpha;
        fg += scale * ci.y * alpha;
        fb += scale * ci.z * alpha;
        scale *= (1.0f - alpha);
    }
    float alpha = ((float)((bg_color & 0xFF000000) >> 24)) / 255.0;
    fr += scale * ((float)((bg_color & 0xFF0000) >> 16)) * alpha;
    fg += scale * ((float)((bg_color & 0xFF00) >> 8)) * alpha;
    fb += scale * ((float)(bg_color & 0xFF)) * alpha;
    scale *= (1.0f - alpha);
    unsigned a;
    if (n < alpha_depth) a = floorf(255 * (1.0f - scale));
    else a = 255;
    unsigned red = floorf(fr / (1.0f - scale));
    unsigned green = floorf(fg / (1.0f - scale));
    unsigned blue = floorf(fb / (1.0f - scale));
    pixels[id] = a << 24 | red << 16 | green << 8 | blue;
}

__global__ void fill_u32_kernel(uint32_t *dst, uint32_t value, uint64_t n) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = value;
}

// transform.cu:10-49 with rotate.h:21-30.  mode 0 translate, 1 rotate, 2 rotate about a point
__global__ void transform_kernel(uint64_t n, float *a, int mode, float phi, float3 axis, float3 v) {
    uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    float3 p = f3(a[3 * id], a[3 * id + 1], a[3 * id + 2]);
    if (mode == 0) {
        p = f3(p.x + v.x, p.y + v.y, p.z + v.z);
    } else {
        if (mode == 2) p = f3(p.x - v.x, p.y - v.y, p.z - v.z);
        float cos_phi = cosf(phi);
        float sin_phi = sinf(phi);
        float pn = p.x * axis.x + p.y * axis.y + p.z * axis.z;
        float3 cr = f3(p.y * axis.z - p.z * axis.y, p.z * axis.x - p.x * axis.z, p.x * axis.y - p.y * axis.x);
        float3 nn = f3(axis.x * pn, axis.y * pn, axis.z * pn);
        float om = 1.0f - cos_phi;
        p = f3(p.x * cos_phi + nn.x * om + cr.x * sin_phi, p.y * cos_phi + nn.y * om + cr.y * sin_phi,
               p.z * cos_phi + nn.z * om + cr.z * sin_phi);
        if (mode == 2) p = f3(p.x + v.x, p.y + v.y, p.z + v.z);
    }
    a[3 * id] = p.x; a[3 * id + 1] = p.y; a[3 * id + 2] = p.z;
}

extern "C" int chroma_rays_destroy(chroma_rays *r) {
    if (!r) return 0;
    cudaSetDevice(r->ctx->device);
    cudaStreamSynchronize(r->ctx->stream);
    void *ptrs[] = {r->d_pos, r->d_dir, r->d_dx, r->d_color, r->d_dxlen};
    for (void *p : ptrs) if (p) cudaFree(p);
    delete r;
    return 0;
}

extern "C" int chroma_rays_create(chroma_ctx *c, uint64_t n, const float *pos, const float *dir,
                                  uint32_t max_alpha_depth, chroma_rays **out) {
    CTX_BIND(c);
    ARG_CHECK(out && n > 0 && pos && dir && max_alpha_depth >= 1 && max_alpha_depth <= 128);
    chroma_rays *r = new chroma_rays();
    r->ctx = c;
    r->n = n;
    r->max_alpha_depth = max_alpha_depth;
    cudaError_t e = cudaSuccess;
    if (e == cudaSuccess) e = cudaMalloc(&r->d_pos, n * 12);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_dir, n * 12);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_dx, n * max_alpha_depth * 4);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_color, n * max_alpha_depth * 16);
    if (e == cudaSuccess) e = cudaMalloc(&r->d_dxlen, n * 4);
    if (e == cudaSuccess) e = cudaMemcpyAsync(r->d_pos, pos, n * 12, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(r->d_dir, dir, n * 12, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(r->d_dxlen, 0, n * 4, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) {
        chroma_set_error("rays_create: %s", cudaGetErrorString(e));
        cudaGetLastError();
        chroma_rays_destroy(r);
        return CHROMA_ERR_CUDA;
    }
    *out = r;
    return 0;
}

extern "C" int chroma_rays_arrays(chroma_rays *r, void **pos, void **dir, void **dx, void **color,
                                  void **dxlen, uint64_t *n) {
    ARG_CHECK(r);
    if (pos) *pos = r->d_pos;
    if (dir) *dir = r->d_dir;
    if (dx) *dx = r->d_dx;
    if (color) *color = r->d_color;
    if (dxlen) *dxlen = r->d_dxlen;
    if (n) *n = r->n;
    return 0;
}

static int rays_transform(chroma_rays *r, float *a, int mode, float phi, const float *axis, const float *v) {
    chroma_ctx *c = r->ctx;
    float3 ax = axis ? make_float3(axis[0], axis[1], axis[2]) : make_float3(0, 0, 0);
    float3 vv = v ? make_float3(v[0], v[1], v[2]) : make_float3(0, 0, 0);
    transform_kernel<<<(unsigned)((r->n + 255) / 256), 256, 0, c->stream>>>(r->n, a, mode, phi, ax, vv);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_rays_translate(chroma_rays *r, const float v[3]) {
    ARG_CHECK(r && v);
    CTX_BIND(r->ctx);
    return rays_transform(r, r->d_pos, 0, 0.0f, nullptr, v);
}

extern "C" int chroma_rays_rotate(chroma_rays *r, float phi, const float axis[3]) {
    ARG_CHECK(r && axis);
    CTX_BIND(r->ctx);
    int rc = rays_transform(r, r->d_pos, 1, phi, axis, nullptr);
    return rc ? rc : rays_transform(r, r->d_dir, 1, phi, axis, nullptr);
}

extern "C" int chroma_rays_rotate_around_point(chroma_rays *r, float phi, const float axis[3], const float point[3]) {
    ARG_CHECK(r && axis && point);
    CTX_BIND(r->ctx);
    int rc = rays_transform(r, r->d_pos, 2, phi, axis, point);
    return rc ? rc : rays_transform(r, r->d_dir, 1, phi, axis, nullptr);
}

extern "C" int chroma_rays_render(chroma_rays *r, chroma_geometry *g, uint32_t alpha_depth, int keep_last_render,
                                  uint32_t bg_color, uint32_t *pixels_device) {
    if (g && g->instanced) {
        chroma_set_error("render needs the flat geometry (colours and the k-nearest list are per flat triangle); "
                         "create the GPUGeometry without instanced=True");
        return CHROMA_ERR_STATE;
    }
    ARG_CHECK(r && g && pixels_device && alpha_depth >= 1);
    chroma_ctx *c = r->ctx;
    CTX_BIND(c);
    if (g->ctx != c) {
        chroma_set_error("rays and geometry belong to different contexts");
        return CHROMA_ERR_ARG;
    }
    if (alpha_depth > r->max_alpha_depth) {   // gpu/render.py:49-50
        chroma_set_error("alpha_depth > max_alpha_depth");
        return CHROMA_ERR_ARG;
    }
    if (!keep_last_render) CU_CHECK(cudaMemsetAsync(r->d_dxlen, 0, r->n * 4, c->stream));
    if (g->wide.ngroups == 0) {
        // no wide tree: either switched off (CHROMA_WIDE=0) or the reference tree is a single leaf,
        // whose triangle the reference never tests (render.cu:66-76 walks the root's zero children)
        uint4 root;
        CU_CHECK(cudaMemcpyAsync(&root, g->d_nodes, sizeof(uint4), cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
        if ((root.w >> 28) != 0) {
            chroma_set_error("render needs the wide search tree (CHROMA_WIDE=0 is set)");
            return CHROMA_ERR_STATE;
        }
        if (keep_last_render) {
            chroma_set_error("keep_last_render on a single-leaf geometry is not supported");
            return CHROMA_ERR_STATE;
        }
        fill_u32_kernel<<<grid_for(r->n, 256, c, 8), 256, 0, c->stream>>>(pixels_device, bg_color, r->n);
        KERNEL_CHECK(c);
        return 0;
    }
    size_t smem = (size_t)WIDE_SSTACK * RENDER_BLOCK * sizeof(uint2) + (size_t)alpha_depth * RENDER_BLOCK * 8;
    CU_CHECK(cudaFuncSetAttribute(render_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    uint32_t *status = c->d_scratch + 2;
    CU_CHECK(cudaMemsetAsync(status, 0, 4, c->stream));
    render_kernel<<<(unsigned)((r->n + RENDER_BLOCK - 1) / RENDER_BLOCK), RENDER_BLOCK, smem, c->stream>>>(
        r->n, r->d_pos, r->d_dir, g->wide, g->d_colors, alpha_depth, pixels_device, r->d_dx, r->d_dxlen,
        r->d_color, bg_color, status);
    KERNEL_CHECK(c);
    CU_CHECK(cudaMemcpyAsync(c->h_scratch + 8, status, 4, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    if (c->h_scratch[8]) {
        chroma_set_error("render: search stack overflow");
        return CHROMA_ERR_OVERFLOW;
    }
    return 0;
}
