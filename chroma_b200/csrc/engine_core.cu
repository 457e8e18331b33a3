// engine_core.cu — context, raw buffers and geometry upload of libchroma_b200.
//
// Replaces chroma/gpu/tools.py:121-168 (context, make_gpu_struct) and
// chroma/gpu/geometry.py:13-259 / chroma/gpu/detector.py:14-39 (GPUGeometry/GPUDetector
// construction).  Instead of the reference's pointer-linked Material/Surface structs the
// optical tables go into one float blob (layout in common.cuh) whose wavelength part is
// staged in shared memory by the kernels; BVH leaves are renumbered into "slots" in leaf
// order and each slot gets one 48-byte packed triangle record, so a leaf test is three
// coalescable 16-byte loads instead of the reference's index load + three dependent
// vertex gathers (chroma/cuda/geometry.h:49-60).
#include <stdarg.h>
#include <stdlib.h>
#include <algorithm>
#include <cub/device/device_scan.cuh>
#include "common.cuh"

static thread_local char g_err[1024] = "";

void chroma_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char *chroma_last_error(void) { return g_err; }
extern "C" const char *chroma_version(void) {
#if CHROMA_AB_KERNELS
    return "chroma_b200 0.2 (sm_100a) +ab";   // variant build with round 1's A/B kernels
#else
    return "chroma_b200 0.2 (sm_100a)";
#endif
}

// ---------------------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------------------
extern "C" int chroma_ctx_create(int device, chroma_ctx **out) {
    ARG_CHECK(out != nullptr);
    int ndev = 0;
    CU_CHECK(cudaGetDeviceCount(&ndev));
    if (device < 0) device = 0;
    if (device >= ndev) {
        chroma_set_error("device %d requested but only %d CUDA device(s) visible", device, ndev);
        return CHROMA_ERR_CUDA;
    }
    CU_CHECK(cudaSetDevice(device));
    chroma_ctx *c = new chroma_ctx();
    c->device = device;
    cudaDeviceProp prop;
    CU_CHECK(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    c->l2_bytes = (size_t)prop.l2CacheSize;
    c->l2_persist_max = (size_t)prop.persistingL2CacheMaxSize;
    c->cc = prop.major * 10 + prop.minor;
    chroma_ctx_read_knobs(c);
    CU_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU_CHECK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    CU_CHECK(cudaEventCreateWithFlags(&c->copy_fork, cudaEventDisableTiming));
    CU_CHECK(cudaMallocHost(&c->h_ready_ring, READY_RING * sizeof(unsigned long long)));
    {   // photon containers and upload staging come from the stream-ordered pool: keep what it
        // has grown to instead of returning it to the driver at every synchronise
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
            uint64_t keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        cudaGetLastError();
    }
    CU_CHECK(cudaEventCreate(&c->ev0));
    CU_CHECK(cudaEventCreate(&c->ev1));
    CU_CHECK(cudaMalloc(&c->d_scratch, 64 * sizeof(uint32_t)));
    CU_CHECK(cudaMemset(c->d_scratch, 0, 64 * sizeof(uint32_t)));
    CU_CHECK(cudaMallocHost(&c->h_scratch, 64 * sizeof(uint32_t)));
    // let kernels keep BVH levels resident in L2 (access-policy window is set per geometry)
    if (c->l2_persist_max > 0)
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, c->l2_persist_max);
    cudaGetLastError();
    *out = c;
    return 0;
}

static int env_int(const char *name, int dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (int)strtol(s, nullptr, 10);
}

// every tuning knob of the context, read in one place (defaults: DESIGN.md section 6 sweeps)
void chroma_ctx_read_knobs(chroma_ctx *c) {
    chroma_ctx::Knobs k;
    k.refill = std::min(32, std::max(1, env_int("CHROMA_REFILL", k.refill)));
    k.iters = std::max(1, env_int("CHROMA_WIDE_ITERS", k.iters));
    k.leaf_thresh = std::max(1, env_int("CHROMA_WIDE_LEAF_THRESH", k.leaf_thresh));
    k.isect_refill = std::max(1, env_int("CHROMA_ISECT_REFILL", k.isect_refill));
    k.isect_iters = std::max(1, env_int("CHROMA_ISECT_ITERS", k.isect_iters));
    k.l2_window_mb = std::max(0, env_int("CHROMA_L2_WINDOW_MB", k.l2_window_mb));
    k.upload_direct = env_int("CHROMA_UPLOAD_DIRECT", k.upload_direct);
    k.prop_v = env_int("CHROMA_PROP_V", k.prop_v);
    k.strict = env_int("CHROMA_STRICT", k.strict);
    k.rng_ool = env_int("CHROMA_RNG_OOL", k.rng_ool);
    k.isect_v = env_int("CHROMA_ISECT_V", k.isect_v);
    k.count_wide = env_int("CHROMA_COUNT_WIDE", k.count_wide);
    k.upload_chunks = std::min(1024, std::max(0, env_int("CHROMA_UPLOAD_CHUNKS", k.upload_chunks)));
    k.upload_threads = std::min(64, std::max(1, env_int("CHROMA_UPLOAD_THREADS", k.upload_threads)));
    k.watchdog_spins = (unsigned)std::max(1, env_int("CHROMA_UPLOAD_WATCHDOG", (int)k.watchdog_spins));
    c->knobs = k;
}
extern "C" int chroma_ctx_reload_knobs(chroma_ctx *c) {
    ARG_CHECK(c);
    chroma_ctx_read_knobs(c);
    return 0;
}

extern "C" int chroma_ctx_destroy(chroma_ctx *c) {
    if (!c) return 0;
    cudaSetDevice(c->device);
    uploader_join(c);
    cudaStreamSynchronize(c->stream);
    if (c->flush_buf) cudaFree(c->flush_buf);
    for (int k = 0; k < c->n_wave_streams; k++) {
        cudaStreamDestroy(c->wave_stream[k]);
        cudaEventDestroy(c->wave_event[k]);
    }
    if (c->wave_fork) cudaEventDestroy(c->wave_fork);
    cudaFree(c->d_scratch);
    cudaFreeHost(c->h_scratch);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaStreamSynchronize(c->copy_stream);
    cudaStreamDestroy(c->copy_stream);
    cudaEventDestroy(c->copy_fork);
    cudaFreeHost(c->h_ready_ring);
    if (c->h_stage_ring) cudaFreeHost(c->h_stage_ring);
    for (cudaEvent_t e : c->stage_slot_done) if (e) cudaEventDestroy(e);
    if (c->d_side) cudaFree(c->d_side);
    cudaStreamDestroy(c->stream);
    delete c;
    return 0;
}

extern "C" int chroma_ctx_synchronize(chroma_ctx *c) {
    CTX_BIND(c);
    if (c->uploader) {      // a borrowed-source upload: its sources are free again when this returns
        const int rc = uploader_join(c);
        CU_CHECK(cudaStreamSynchronize(c->copy_stream));
        if (rc) return rc;
    }
    if (c->status_pending) {
        // status word of the last propagate launch that was not asked for it (engine_wide.cu)
        CU_CHECK(cudaMemcpyAsync(c->h_scratch + 7, c->d_scratch + 7, 4, cudaMemcpyDeviceToHost, c->stream));
        CU_CHECK(cudaStreamSynchronize(c->stream));
        c->status_pending = false;
        c->sticky_status |= c->h_scratch[7];
    } else {
        CU_CHECK(cudaStreamSynchronize(c->stream));
    }
    if (c->sticky_status) {
        const uint32_t s = c->sticky_status;
        c->sticky_status = 0;
        return chroma_status_to_error(s);
    }
    return 0;
}

extern "C" int chroma_ctx_stream(chroma_ctx *c, void **stream) {
    ARG_CHECK(c && stream);
    *stream = (void *)c->stream;
    return 0;
}

extern "C" int chroma_ctx_info(chroma_ctx *c, uint64_t out[6]) {
    CTX_BIND(c);
    size_t fr = 0, tot = 0;
    CU_CHECK(cudaMemGetInfo(&fr, &tot));
    out[0] = fr; out[1] = tot; out[2] = (uint64_t)c->sm_count; out[3] = c->l2_bytes;
    out[4] = c->l2_persist_max; out[5] = (uint64_t)c->cc;
    return 0;
}

extern "C" int chroma_ctx_launch_count(chroma_ctx *c, uint64_t *out) {
    ARG_CHECK(c && out);
    *out = c->launches;
    return 0;
}

extern "C" int chroma_ctx_timer_begin(chroma_ctx *c) {
    CTX_BIND(c);
    CU_CHECK(cudaEventRecord(c->ev0, c->stream));
    return 0;
}

extern "C" int chroma_ctx_timer_end(chroma_ctx *c, float *ms) {
    CTX_BIND(c);
    ARG_CHECK(ms != nullptr);
    CU_CHECK(cudaEventRecord(c->ev1, c->stream));
    CU_CHECK(cudaEventSynchronize(c->ev1));
    CU_CHECK(cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return 0;
}

extern "C" int chroma_ctx_flush_l2(chroma_ctx *c) {
    CTX_BIND(c);
    if (!c->flush_buf) {
        c->flush_bytes = c->l2_bytes * 2 + (64u << 20);
        CU_CHECK(cudaMalloc(&c->flush_buf, c->flush_bytes));
    }
    CU_CHECK(cudaMemsetAsync(c->flush_buf, 0x5a, c->flush_bytes, c->stream));
    return 0;
}

// ---------------------------------------------------------------------------------------
// raw buffers
// ---------------------------------------------------------------------------------------
extern "C" int chroma_malloc(chroma_ctx *c, size_t bytes, void **dptr) {
    CTX_BIND(c);
    ARG_CHECK(dptr != nullptr);
    if (bytes == 0) bytes = 16;  // forbid zero-size, as the reference does for extra_nodes
    // stream-ordered pool (release threshold set in chroma_ctx_create): cudaMalloc / cudaFree cost
    // a device-wide synchronisation and, next to tens of GB of geometry, stalls of 10 ms to 1 s
    // (tools/bm_probe.py, s19 logs): a GPUArray-style temporary per call must not pay that
    CU_CHECK(cudaMallocAsync(dptr, bytes, c->stream));
    return 0;
}
extern "C" int chroma_free(chroma_ctx *c, void *dptr) {
    CTX_BIND(c);
    if (dptr) CU_CHECK(cudaFreeAsync(dptr, c->stream));   // after everything queued on the stream so far
    return 0;
}
extern "C" int chroma_memcpy_h2d(chroma_ctx *c, void *dst, const void *src, size_t bytes) {
    CTX_BIND(c);
    if (bytes == 0) return 0;
    CU_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c->stream));
    // pageable sources are staged by the runtime before return; pinned ones need the sync
    CU_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}
extern "C" int chroma_memcpy_d2h(chroma_ctx *c, void *dst, const void *src, size_t bytes) {
    CTX_BIND(c);
    if (bytes == 0) return 0;
    CU_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}
extern "C" int chroma_memcpy_d2d(chroma_ctx *c, void *dst, const void *src, size_t bytes) {
    CTX_BIND(c);
    if (bytes == 0) return 0;
    CU_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, c->stream));
    return 0;
}

__global__ void fill32_kernel(uint32_t *dst, uint32_t value, size_t count) {
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        dst[i] = value;
}

extern "C" int chroma_memset32(chroma_ctx *c, void *dst, uint32_t value, size_t count) {
    CTX_BIND(c);
    if (count == 0) return 0;
    fill32_kernel<<<grid_for(count, 256, c, 8), 256, 0, c->stream>>>((uint32_t *)dst, value, count);
    KERNEL_CHECK(c);
    return 0;
}

extern "C" int chroma_host_alloc(size_t bytes, void **hptr) {
    ARG_CHECK(hptr != nullptr);
    CU_CHECK(cudaMallocHost(hptr, bytes ? bytes : 16));
    return 0;
}
extern "C" int chroma_host_free(void *hptr) {
    if (hptr) CU_CHECK(cudaFreeHost(hptr));
    return 0;
}

// ---------------------------------------------------------------------------------------
// geometry build kernels
// ---------------------------------------------------------------------------------------
// one sweep of "children of reachable inner nodes are reachable"
__global__ void mark_children_kernel(const uint4 *nodes, uint64_t nnodes, uint32_t *reach,
                                     uint32_t *changed) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnodes; i += stride) {
        if (!reach[i]) continue;
        uint32_t w = nodes[i].w;
        uint32_t nchild = w >> 28, first = w & 0x0FFFFFFFu;
        for (uint32_t c = first; c < first + nchild && c < nnodes; c++) {
            if (!reach[c]) {
                reach[c] = 1;
                *changed = 1;
            }
        }
    }
}

// leaf flags as a bit mask (kept with the geometry: 1 bit per node) and the reference-format node array
// rebuilt from the engine's copy: they differ only in the leaves' w (slot vs triangle id)
__global__ void leaf_bits_kernel(const uint32_t *flag, uint64_t nnodes, uint32_t *bits) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnodes; i += stride) {
        const unsigned m = __ballot_sync(__activemask(), flag[i] != 0);
        if ((threadIdx.x & 31) == 0) bits[i >> 5] = m;   // block size and stride are multiples of 32: a warp covers one aligned word
    }
}
__global__ void ref_nodes_kernel(const uint4 *nodes, uint64_t nnodes, const uint32_t *bits, const float4 *tris, uint4 *out) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnodes; i += stride) {
        uint4 n = nodes[i];
        if ((bits[i >> 5] >> (i & 31)) & 1u) n.w = __float_as_uint(tris[(size_t)n.w * 3 + 2].z);
        out[i] = n;
    }
}

__global__ void leaf_flag_kernel(const uint4 *nodes, uint64_t nnodes, const uint32_t *reach,
                                 uint32_t *flag) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnodes; i += stride)
        flag[i] = (reach[i] && (nodes[i].w >> 28) == 0) ? 1u : 0u;
}

// write engine nodes (leaf w = slot) and the packed triangle records
__global__ void build_engine_kernel(const uint4 *nodes, uint64_t nnodes, const uint32_t *flag,
                                    const uint32_t *slot_of, const float *vertices,
                                    const uint32_t *triangles, const uint32_t *codes,
                                    const uint32_t *solid_id, uint64_t ntriangles,
                                    uint64_t nvertices, uint4 *out_nodes, float4 *tris,
                                    int32_t *tri_slot, uint32_t *bad) {
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnodes; i += stride) {
        uint4 n = nodes[i];
        if (flag[i]) {
            uint32_t tri = n.w;  // nchild == 0 so w is the triangle id
            uint32_t slot = slot_of[i];
            if (tri >= ntriangles) {
                atomicAdd(bad, 1u);
                n.w = 0;  // keep traversal in bounds
            } else {
                uint32_t i0 = triangles[3 * (size_t)tri], i1 = triangles[3 * (size_t)tri + 1],
                         i2 = triangles[3 * (size_t)tri + 2];
                if (i0 >= nvertices || i1 >= nvertices || i2 >= nvertices) {
                    atomicAdd(bad, 1u);
                    i0 = i1 = i2 = 0;
                }
                const float *v0 = vertices + 3 * (size_t)i0, *v1 = vertices + 3 * (size_t)i1,
                            *v2 = vertices + 3 * (size_t)i2;
                float4 *rec = tris + 3 * (size_t)slot;
                rec[0] = make_float4(v0[0], v0[1], v0[2], v1[0]);
                rec[1] = make_float4(v1[1], v1[2], v2[0], v2[1]);
                rec[2] = make_float4(v2[2], __uint_as_float(codes[tri]), __uint_as_float(tri),
                                     __uint_as_float(solid_id ? solid_id[tri] : 0u));
                tri_slot[tri] = (int32_t)slot;
                n.w = slot;
            }
        }
        out_nodes[i] = n;
    }
}

template <typename T>
static int upload(chroma_ctx *c, T **dptr, const T *src, size_t count, uint64_t *bytes) {
    size_t nb = (count ? count : 1) * sizeof(T);
    CU_CHECK(cudaMalloc((void **)dptr, nb));
    if (bytes) *bytes += nb;
    if (src && count)
        CU_CHECK(cudaMemcpyAsync(*dptr, src, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    else
        CU_CHECK(cudaMemsetAsync(*dptr, 0, nb, c->stream));
    return 0;
}

#define TRY(x)                                                                             \
    do {                                                                                   \
        int _r = (x);                                                                      \
        if (_r) return _r;                                                                 \
    } while (0)

static uint32_t env_u32(const char *name, uint32_t dflt) {
    const char *s = getenv(name);
    if (!s || !*s) return dflt;
    return (uint32_t)strtoul(s, nullptr, 10);
}

int chroma_build_optics(chroma_ctx *c, chroma_geometry *g, const chroma_geometry_desc *d) {
    OpticsView &o = g->view.opt;
    const uint32_t nw = d->nwavelengths, nt = d->ntimes;
    ARG_CHECK(nw >= 2);
    // material / surface indices travel as signed 8-bit fields of the triangle's code word
    // (photon.h:68-75 convert: 0x80..0xFF sign-extend, 0xFF = none), so at most 128 of each
    if (d->nmaterials > 128 || d->nsurfaces > 128) {
        chroma_set_error("%u materials / %u surfaces: at most 128 of each fit the 8-bit signed index of a "
                         "triangle's material code (chroma/cuda/photon.h:68-75)", d->nmaterials, d->nsurfaces);
        return CHROMA_ERR_ARG;
    }
    ARG_CHECK(d->nmaterials >= 1 && d->material_props);
    ARG_CHECK(d->ncomp_total == 0 ||
              (d->comp_reemission_prob && d->comp_reemission_wvl_cdf &&
               d->comp_absorption_length && d->comp_reemission_time_cdf && nt >= 2));
    o.nw = nw; o.nt = nt;
    o.w0 = d->wavelength_start; o.dw = d->wavelength_step;
    o.t0 = d->time_start; o.dt = d->time_step;
    o.nmat = d->nmaterials; o.nsurf = d->nsurfaces;
    g->physics_heavy = d->ncomp_total > 0;
    for (uint32_t sidx = 0; sidx < d->nsurfaces && d->surface_model; sidx++)
        if (d->surface_model[sidx] != 0 && (!d->surface_present || d->surface_present[sidx])) g->physics_heavy = true;

    std::vector<float> blob;
    o.mat_off = 0;
    blob.insert(blob.end(), d->material_props, d->material_props + (size_t)d->nmaterials * 3 * nw);
    o.surf_off = (uint32_t)blob.size();
    if (d->nsurfaces) {
        ARG_CHECK(d->surface_props && d->surface_model && d->surface_transmissive &&
                  d->surface_thickness);
        blob.insert(blob.end(), d->surface_props, d->surface_props + (size_t)d->nsurfaces * 8 * nw);
    }
    o.comp_off = (uint32_t)blob.size();
    for (uint32_t cidx = 0; cidx < d->ncomp_total; cidx++) {
        const float *a = d->comp_reemission_prob + (size_t)cidx * nw;
        const float *b = d->comp_reemission_wvl_cdf + (size_t)cidx * nw;
        const float *e = d->comp_absorption_length + (size_t)cidx * nw;
        blob.insert(blob.end(), a, a + nw);
        blob.insert(blob.end(), b, b + nw);
        blob.insert(blob.end(), e, e + nw);
    }
    const uint32_t stageable = (uint32_t)blob.size();  // everything so far is wavelength tables
    // dichroic blocks
    std::vector<uint4> meta(d->nsurfaces ? d->nsurfaces : 1, make_uint4(0, 0, 0xFFFFFFFFu, 0));
    size_t acc = 0;
    for (uint32_t s = 0; s < d->nsurfaces; s++) {
        meta[s].x = d->surface_model[s];
        meta[s].y = d->surface_transmissive[s];
        uint32_t na = d->dichroic_nangles ? d->dichroic_nangles[s] : 0;
        if (na) {
            ARG_CHECK(d->dichroic_angles && d->dichroic_reflect && d->dichroic_transmit);
            meta[s].z = (uint32_t)blob.size();
            meta[s].w = na;
            blob.insert(blob.end(), d->dichroic_angles + acc, d->dichroic_angles + acc + na);
            blob.insert(blob.end(), d->dichroic_reflect + acc * nw, d->dichroic_reflect + (acc + na) * nw);
            blob.insert(blob.end(), d->dichroic_transmit + acc * nw, d->dichroic_transmit + (acc + na) * nw);
            acc += na;
        } else if (d->surface_model[s] == CHROMA_SURFACE_DICHROIC && d->surface_present &&
                   d->surface_present[s]) {
            chroma_set_error("surface %u uses the dichroic model but has no dichroic_props", s);
            return CHROMA_ERR_ARG;
        }
    }
    o.tcdf_off = (uint32_t)blob.size();
    if (d->ncomp_total)
        blob.insert(blob.end(), d->comp_reemission_time_cdf,
                    d->comp_reemission_time_cdf + (size_t)d->ncomp_total * nt);
    blob.push_back(0.0f);  // interp_property may touch fp[n] with zero weight (geometry.h:74)
    blob.push_back(0.0f);

    std::vector<uint32_t> ncomp(d->nmaterials), cfirst(d->nmaterials);
    uint32_t run = 0;
    for (uint32_t m = 0; m < d->nmaterials; m++) {
        ncomp[m] = d->material_num_comp ? d->material_num_comp[m] : 0;
        cfirst[m] = run;
        run += ncomp[m];
    }
    ARG_CHECK(run == d->ncomp_total);

    TRY(upload(c, &g->d_optics, blob.data(), blob.size(), &g->bytes));
    TRY(upload(c, &g->d_mat_ncomp, ncomp.data(), ncomp.size(), &g->bytes));
    TRY(upload(c, &g->d_mat_comp_first, cfirst.data(), cfirst.size(), &g->bytes));
    TRY(upload(c, &g->d_surf_meta, meta.data(), meta.size(), &g->bytes));
    TRY(upload(c, &g->d_surf_thickness, d->surface_thickness, (size_t)d->nsurfaces, &g->bytes));
    CU_CHECK(cudaStreamSynchronize(c->stream));  // host vectors go out of scope
    o.blob = g->d_optics;
    o.mat_ncomp = g->d_mat_ncomp;
    o.mat_comp_first = g->d_mat_comp_first;
    o.surf_meta = g->d_surf_meta;
    o.surf_thickness = g->d_surf_thickness;
    // stage the wavelength tables in shared memory when they fit the budget
    // r01h/r01v: reading the tables through L1 costs nothing measurable and the shared memory is
    // worth more as search stack (4 CTAs/SM x 20 entries/thread); staging stays available
    uint32_t budget = env_u32("CHROMA_OPTICS_SMEM_FLOATS", 0);
    o.smem_floats = stageable <= budget ? stageable : 0;
    return 0;
}

extern "C" int chroma_geometry_create(chroma_ctx *c, const chroma_geometry_desc *d,
                                      chroma_geometry **out) {
    CTX_BIND(c);
    ARG_CHECK(d && out);
    ARG_CHECK(d->vertices && d->triangles && d->material_codes);
    ARG_CHECK(d->ntriangles > 0 && d->ntriangles < (1ull << 28));
    ARG_CHECK(d->nvertices > 0 && d->nvertices < (1ull << 32));
    ARG_CHECK((d->nodes && d->nnodes > 0) || d->bvh);
    chroma_geometry *g = new chroma_geometry();
    g->ctx = c;
    g->ntriangles = d->ntriangles;
    g->nvertices = d->nvertices;

    int rc = 0;
    uint32_t *d_reach = nullptr, *d_flag = nullptr, *d_scan = nullptr;
    void *d_temp = nullptr;
    do {
        if ((rc = upload(c, &g->d_vertices, d->vertices, (size_t)d->nvertices * 3, &g->bytes))) break;
        if ((rc = upload(c, &g->d_triangles, d->triangles, (size_t)d->ntriangles * 3, &g->bytes))) break;
        if ((rc = upload(c, &g->d_material_codes, d->material_codes, (size_t)d->ntriangles, &g->bytes))) break;
        if ((rc = upload(c, &g->d_colors, d->colors, (size_t)d->ntriangles, &g->bytes))) break;
        if ((rc = upload(c, &g->d_solid_id, d->solid_id, (size_t)d->ntriangles, &g->bytes))) break;
        if (d->bvh) {
            g->nnodes = d->bvh->nnodes;
            size_t nb = g->nnodes * sizeof(uint4);
            if (cudaMalloc(&g->d_nodes_ref, nb) != cudaSuccess) { rc = CHROMA_ERR_CUDA; chroma_set_error("cudaMalloc nodes"); break; }
            g->bytes += nb;
            cudaMemcpyAsync(g->d_nodes_ref, d->bvh->d_nodes, nb, cudaMemcpyDeviceToDevice, c->stream);
            g->view.world_origin = make_float3(d->bvh->world_origin[0], d->bvh->world_origin[1], d->bvh->world_origin[2]);
            g->view.world_scale = d->bvh->world_scale;
        } else {
            g->nnodes = d->nnodes;
            if ((rc = upload(c, &g->d_nodes_ref, (const uint4 *)d->nodes, (size_t)d->nnodes, &g->bytes))) break;
            g->view.world_origin = make_float3(d->world_origin[0], d->world_origin[1], d->world_origin[2]);
            g->view.world_scale = d->world_scale;
        }
        const uint64_t nn = g->nnodes;
        if (nn >= (1ull << 28)) { chroma_set_error("too many BVH nodes"); rc = CHROMA_ERR_ARG; break; }
        // reachability fixed point from the root
        if (cudaMalloc(&d_reach, nn * 4) != cudaSuccess || cudaMalloc(&d_flag, nn * 4) != cudaSuccess ||
            cudaMalloc(&d_scan, nn * 4) != cudaSuccess) {
            chroma_set_error("cudaMalloc (geometry scratch): %s", cudaGetErrorString(cudaGetLastError()));
            rc = CHROMA_ERR_CUDA; break;
        }
        cudaMemsetAsync(d_reach, 0, nn * 4, c->stream);
        uint32_t one = 1;
        cudaMemcpyAsync(d_reach, &one, 4, cudaMemcpyHostToDevice, c->stream);
        unsigned grid = grid_for(nn, 256, c, 8);
        int iter = 0;
        for (; iter < 256; iter++) {
            cudaMemsetAsync(c->d_scratch, 0, 4, c->stream);
            mark_children_kernel<<<grid, 256, 0, c->stream>>>(g->d_nodes_ref, nn, d_reach, c->d_scratch);
            c->launches++;
            cudaMemcpyAsync(c->h_scratch, c->d_scratch, 4, cudaMemcpyDeviceToHost, c->stream);
            if (cudaStreamSynchronize(c->stream) != cudaSuccess) break;
            if (c->h_scratch[0] == 0) break;
        }
        if (cudaGetLastError() != cudaSuccess || iter == 256) {
            chroma_set_error("BVH reachability sweep failed (cyclic or corrupt node array?)");
            rc = CHROMA_ERR_ARG; break;
        }
        leaf_flag_kernel<<<grid, 256, 0, c->stream>>>(g->d_nodes_ref, nn, d_reach, d_flag);
        c->launches++;
        size_t temp_bytes = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, d_flag, d_scan, (int)nn, c->stream);
        if (cudaMalloc(&d_temp, temp_bytes ? temp_bytes : 16) != cudaSuccess) { rc = CHROMA_ERR_CUDA; chroma_set_error("cudaMalloc scan temp"); break; }
        cub::DeviceScan::ExclusiveSum(d_temp, temp_bytes, d_flag, d_scan, (int)nn, c->stream);
        c->launches++;
        uint32_t last_flag = 0, last_scan = 0;
        cudaMemcpyAsync(&last_flag, d_flag + nn - 1, 4, cudaMemcpyDeviceToHost, c->stream);
        cudaMemcpyAsync(&last_scan, d_scan + nn - 1, 4, cudaMemcpyDeviceToHost, c->stream);
        if (cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = CHROMA_ERR_CUDA; chroma_set_error("scan: %s", cudaGetErrorString(cudaGetLastError())); break; }
        g->nslots = (uint64_t)last_flag + last_scan;
        // engine arrays
        size_t nb_nodes = nn * sizeof(uint4), nb_tris = (g->nslots ? g->nslots : 1) * 3 * sizeof(float4);
        if (cudaMalloc(&g->d_nodes, nb_nodes) != cudaSuccess || cudaMalloc(&g->d_tris, nb_tris) != cudaSuccess ||
            cudaMalloc(&g->d_tri_slot, g->ntriangles * 4) != cudaSuccess) {
            chroma_set_error("cudaMalloc (engine arrays): %s", cudaGetErrorString(cudaGetLastError()));
            rc = CHROMA_ERR_CUDA; break;
        }
        g->bytes += nb_nodes + nb_tris + g->ntriangles * 4;
        cudaMemsetAsync(g->d_tri_slot, 0xFF, g->ntriangles * 4, c->stream);
        cudaMemsetAsync(c->d_scratch, 0, 4, c->stream);
        build_engine_kernel<<<grid, 256, 0, c->stream>>>(
            g->d_nodes_ref, nn, d_flag, d_scan, g->d_vertices, g->d_triangles, g->d_material_codes,
            d->solid_id ? g->d_solid_id : nullptr, g->ntriangles, g->nvertices, g->d_nodes, g->d_tris,
            g->d_tri_slot, c->d_scratch);
        c->launches++;
        cudaMemcpyAsync(c->h_scratch, c->d_scratch, 4, cudaMemcpyDeviceToHost, c->stream);
        if (cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = CHROMA_ERR_CUDA; chroma_set_error("build_engine: %s", cudaGetErrorString(cudaGetLastError())); break; }
        if (c->h_scratch[0] != 0) {
            chroma_set_error("%u BVH leaves refer to triangles/vertices out of range", c->h_scratch[0]);
            rc = CHROMA_ERR_ARG; break;
        }
        if ((rc = chroma_build_optics(c, g, d))) break;
        g->view.nodes = g->d_nodes;
        g->view.tris = g->d_tris;
        g->view.tri_slot = g->d_tri_slot;
        g->view.nnodes = (uint32_t)nn;
        g->view.nslots = (uint32_t)g->nslots;
        g->view.ntriangles = (uint32_t)g->ntriangles;
        uint32_t top = env_u32("CHROMA_TOP_NODES", 512);
        g->view.top_nodes = (uint32_t)(nn < top ? nn : top);
        // the engine's own search structure over the same triangles (engine_wide.cu)
        if ((rc = chroma_wide_build(g, d_reach, d_flag, d_scan))) break;
        g->tree_kind = g->wide.ngroups ? 1 : 0;
        // The reference-format copy of the nodes (3.9 GB at C3) is needed by nothing once the engine's
        // arrays exist; it is rebuilt from them if GPUGeometry.nodes is ever read (chroma_geometry_array 0)
        if (!env_u32("CHROMA_KEEP_REF_NODES", 0)) {
            const uint64_t nwords = (nn + 31) / 32 + 1;
            if (cudaMalloc(&g->d_leaf_bits, nwords * 4) != cudaSuccess) { rc = CHROMA_ERR_CUDA; chroma_set_error("cudaMalloc leaf bits"); break; }
            // grid of whole warps over consecutive nodes: thread i handles node i, i + stride, ... with
            // stride a multiple of 32, so every warp iteration covers one aligned 32-node word
            leaf_bits_kernel<<<grid, 256, 0, c->stream>>>(d_flag, nn, g->d_leaf_bits);
            c->launches++;
            if (cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = CHROMA_ERR_CUDA; chroma_set_error("leaf bits: %s", cudaGetErrorString(cudaGetLastError())); break; }
            cudaFree(g->d_nodes_ref);
            g->d_nodes_ref = nullptr;
            g->bytes -= nn * sizeof(uint4);
            g->bytes += nwords * 4;
        }
    } while (0);
    if (d_reach) cudaFree(d_reach);
    if (d_flag) cudaFree(d_flag);
    if (d_scan) cudaFree(d_scan);
    if (d_temp) cudaFree(d_temp);
    if (rc) {
        chroma_geometry_destroy(g);
        return rc;
    }
    *out = g;
    return 0;
}

extern "C" int chroma_geometry_destroy(chroma_geometry *g) {
    if (!g) return 0;
    cudaSetDevice(g->ctx->device);
    cudaStreamSynchronize(g->ctx->stream);
    void *ptrs[] = {g->d_nodes_ref, g->d_vertices, g->d_triangles, g->d_material_codes, g->d_colors,
                    g->d_solid_id, g->d_nodes, g->d_tris, g->d_tri_slot, g->d_optics, g->d_mat_ncomp,
                    g->d_mat_comp_first, g->d_surf_meta, g->d_surf_thickness, g->d_solid_to_channel,
                    g->d_time_cdf_x, g->d_time_cdf_y, g->d_charge_cdf_x, g->d_charge_cdf_y,
                    g->d_wgroups, g->d_wtris, g->d_wrank, g->d_wtri_slot,
                    g->d_inst, g->d_tri_base, g->d_tmpl_off_of_orig, g->d_tmpl_colors,
                    g->d_tmpl_vert, g->d_tmpl_tri, g->d_tmpl_mcode, g->d_tmpl_nvert, g->d_vert_base,
                    g->d_leaf_bits, g->d_tmpl_info, g->d_placements, g->d_flat[1], g->d_flat[2], g->d_flat[3], g->d_flat[4],
                    g->d_flat[5]};
    // an access-policy window that still points into this geometry is withdrawn before its memory goes
    if (g->l2_window_stream) {
        cudaStreamAttrValue av{};
        av.accessPolicyWindow.num_bytes = 0;
        av.accessPolicyWindow.hitProp = cudaAccessPropertyNormal;
        av.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
        cudaStreamSetAttribute(g->l2_window_stream, cudaStreamAttributeAccessPolicyWindow, &av);
        cudaCtxResetPersistingL2Cache();
        cudaGetLastError();
    }
    for (void *p : ptrs)
        if (p) cudaFree(p);
    delete g;
    return 0;
}

extern "C" int chroma_geometry_array(chroma_geometry *g, int which, void **dptr, uint64_t *count) {
    ARG_CHECK(g && dptr && count);
    if (g->instanced) {
        CTX_BIND(g->ctx);
        return chroma_inst_flat_array(g, which, dptr, count);
    }
    if (which == 0 && !g->d_nodes_ref && g->d_leaf_bits) {
        chroma_ctx *c = g->ctx;
        CTX_BIND(c);
        CU_CHECK(cudaMalloc(&g->d_nodes_ref, g->nnodes * sizeof(uint4)));
        g->bytes += g->nnodes * sizeof(uint4);
        ref_nodes_kernel<<<grid_for(g->nnodes, 256, c, 8), 256, 0, c->stream>>>(g->d_nodes, g->nnodes, g->d_leaf_bits, g->d_tris,
                                                                                  g->d_nodes_ref);
        KERNEL_CHECK(c);
    }
    if (which >= 1 && which <= 5 && which != 4 && g->mesh_released) {
        chroma_set_error("the mesh arrays of this geometry were released (chroma_geometry_release_mesh): the caller "
                         "holds them on the host");
        return CHROMA_ERR_STATE;
    }
    switch (which) {
        case 0: *dptr = g->d_nodes_ref; *count = g->nnodes; break;
        case 1: *dptr = g->d_vertices; *count = g->nvertices; break;
        case 2: *dptr = g->d_triangles; *count = g->ntriangles; break;
        case 3: *dptr = g->d_material_codes; *count = g->ntriangles; break;
        case 4: *dptr = g->d_colors; *count = g->ntriangles; break;
        case 5: *dptr = g->d_solid_id; *count = g->ntriangles; break;
        default: chroma_set_error("unknown geometry array %d", which); return CHROMA_ERR_ARG;
    }
    return 0;
}

extern "C" int chroma_geometry_release_mesh(chroma_geometry *g) {
    ARG_CHECK(g);
    if (g->instanced || g->mesh_released) return 0;
    chroma_ctx *c = g->ctx;
    CTX_BIND(c);
    CU_CHECK(cudaStreamSynchronize(c->stream));
    void **ptrs[] = {(void **)&g->d_vertices, (void **)&g->d_triangles, (void **)&g->d_material_codes, (void **)&g->d_solid_id};
    const uint64_t sizes[] = {g->nvertices * 12, g->ntriangles * 12, g->ntriangles * 4, g->ntriangles * 4};
    for (int k = 0; k < 4; k++)
        if (*ptrs[k]) { cudaFree(*ptrs[k]); *ptrs[k] = nullptr; g->bytes -= sizes[k]; }
    g->mesh_released = true;
    return 0;
}

extern "C" int chroma_geometry_info(chroma_geometry *g, uint64_t out[8]) {
    ARG_CHECK(g && out);
    out[0] = g->nnodes; out[1] = g->ntriangles; out[2] = g->nslots; out[3] = g->bytes;
    out[4] = g->view.opt.smem_floats;
    out[5] = g->wide.ngroups; out[6] = g->wide_levels; out[7] = g->wide_leaves;
    return 0;
}

extern "C" int chroma_geometry_tree(chroma_geometry *g, uint64_t out[8]) {
    ARG_CHECK(g && out);
    out[0] = (uint64_t)g->tree_kind;
    out[1] = g->instanced ? g->nsolids : 0;
    out[2] = g->ntemplates;
    out[3] = g->nslots;
    out[4] = g->instanced ? g->inst_top_groups : g->wide.ngroups;
    out[5] = g->wide.ngroups;
    out[6] = (uint64_t)g->wide.ngroups * 128 + g->nslots * (g->instanced ? 48 : 52) +
             (g->instanced ? g->nsolids * 64 : 0);
    out[7] = g->instanced ? g->inst_top_leaves : 0;
    return 0;
}

extern "C" int chroma_detector_attach(chroma_geometry *g, const int32_t *solid_to_channel,
                                      uint64_t nsolids, uint32_t nchannels, const float *time_cdf_x,
                                      const float *time_cdf_y, uint32_t time_cdf_len,
                                      const float *charge_cdf_x, const float *charge_cdf_y,
                                      uint32_t charge_cdf_len) {
    ARG_CHECK(g && solid_to_channel && nsolids > 0);
    ARG_CHECK(time_cdf_x && time_cdf_y && time_cdf_len >= 2);
    ARG_CHECK(charge_cdf_x && charge_cdf_y && charge_cdf_len >= 2);
    chroma_ctx *c = g->ctx;
    CTX_BIND(c);
    if (g->has_detector) {
        chroma_set_error("detector already attached");
        return CHROMA_ERR_STATE;
    }
    TRY(upload(c, &g->d_solid_to_channel, solid_to_channel, nsolids, &g->bytes));
    TRY(upload(c, &g->d_time_cdf_x, time_cdf_x, time_cdf_len, &g->bytes));
    TRY(upload(c, &g->d_time_cdf_y, time_cdf_y, time_cdf_len, &g->bytes));
    TRY(upload(c, &g->d_charge_cdf_x, charge_cdf_x, charge_cdf_len, &g->bytes));
    TRY(upload(c, &g->d_charge_cdf_y, charge_cdf_y, charge_cdf_len, &g->bytes));
    CU_CHECK(cudaStreamSynchronize(c->stream));
    g->nsolids = nsolids;
    g->det.solid_to_channel = g->d_solid_to_channel;
    g->det.time_cdf_x = g->d_time_cdf_x; g->det.time_cdf_y = g->d_time_cdf_y;
    g->det.charge_cdf_x = g->d_charge_cdf_x; g->det.charge_cdf_y = g->d_charge_cdf_y;
    g->det.nchannels = (int)nchannels;
    g->det.time_cdf_len = (int)time_cdf_len;
    g->det.charge_cdf_len = (int)charge_cdf_len;
    // chroma/gpu/detector.py:39: charge_unit = charge_cdf_x[-1] / 2**16
    g->det.charge_unit = charge_cdf_x[charge_cdf_len - 1] / 65536.0f;
    g->has_detector = true;
    return 0;
}
