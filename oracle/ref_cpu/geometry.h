/* oracle/ref_cpu/geometry.h — TEST INFRASTRUCTURE ONLY.  Included after the reference's headers. */
#ifndef CHROMA_REF_CPU_GEOMETRY_H
#define CHROMA_REF_CPU_GEOMETRY_H
#include "chroma_oracle.h"   // oracle_geometry: the flat description both checkers take

namespace {

// The reference's pointer-linked device structs (geometry_types.h), assembled from the flat tables
// the way chroma/gpu/geometry.py:13-259 assembles them on the device.
struct RefGeometry {
    Geometry g;
    std::vector<Material> materials;
    std::vector<Material *> material_ptrs;
    std::vector<std::vector<float *>> comp_ptrs;   // per material: 4 lists of per-component pointers
    std::vector<Surface> surfaces;
    std::vector<Surface *> surface_ptrs;
    std::vector<DichroicProps> dichroic;
    std::vector<std::vector<float *>> dich_ptrs;
#ifdef REF_CPU_WITH_DETECTOR
    Detector det;
#endif

    explicit RefGeometry(const oracle_geometry *o) {
        const uint32_t nw = o->nwavelengths;
        memset(&g, 0, sizeof(g));
        g.vertices = (float3 *)o->vertices;
        g.triangles = (uint3 *)o->triangles;
        g.material_codes = (unsigned int *)o->material_codes;
        g.primary_nodes = (uint4 *)o->nodes;
        g.extra_nodes = nullptr;
        g.nprimary_nodes = (int)o->nnodes;
        g.world_origin = make_float3(o->world_origin[0], o->world_origin[1], o->world_origin[2]);
        g.world_scale = o->world_scale;
        materials.resize(o->nmaterials);
        comp_ptrs.resize((size_t)o->nmaterials * 4);
        uint32_t c0 = 0;
        for (uint32_t m = 0; m < o->nmaterials; m++) {
            Material &M = materials[m];
            float *base = (float *)o->material_props + (size_t)m * 3 * nw;
            M.refractive_index = base; M.absorption_length = base + nw; M.scattering_length = base + 2 * nw;
            const uint32_t nc = o->material_num_comp ? o->material_num_comp[m] : 0;
            for (uint32_t c = 0; c < nc; c++) {
                comp_ptrs[m * 4 + 0].push_back((float *)o->comp_reemission_prob + (size_t)(c0 + c) * nw);
                comp_ptrs[m * 4 + 1].push_back((float *)o->comp_reemission_wvl_cdf + (size_t)(c0 + c) * nw);
                comp_ptrs[m * 4 + 2].push_back((float *)o->comp_reemission_time_cdf + (size_t)(c0 + c) * o->ntimes);
                comp_ptrs[m * 4 + 3].push_back((float *)o->comp_absorption_length + (size_t)(c0 + c) * nw);
            }
            c0 += nc;
            M.comp_reemission_prob = comp_ptrs[m * 4 + 0].data();
            M.comp_reemission_wvl_cdf = comp_ptrs[m * 4 + 1].data();
            M.comp_reemission_time_cdf = comp_ptrs[m * 4 + 2].data();
            M.comp_absorption_length = comp_ptrs[m * 4 + 3].data();
            M.num_comp = nc;
            M.wavelength_n = nw; M.wavelength_step = o->wavelength_step; M.wavelength_start = o->wavelength_start;
            M.time_n = o->ntimes; M.time_step = o->time_step; M.time_start = o->time_start;
        }
        for (auto &M : materials) material_ptrs.push_back(&M);
        g.materials = material_ptrs.data();
        surfaces.resize(o->nsurfaces);
        dichroic.resize(o->nsurfaces);
        dich_ptrs.resize((size_t)o->nsurfaces * 2);
        uint64_t a0 = 0;
        for (uint32_t s = 0; s < o->nsurfaces; s++) {
            Surface &S = surfaces[s];
            memset(&S, 0, sizeof(S));
            float *base = (float *)o->surface_props + (size_t)s * 8 * nw;
            S.detect = base; S.absorb = base + nw; S.reemit = base + 2 * nw; S.reflect_diffuse = base + 3 * nw;
            S.reflect_specular = base + 4 * nw; S.eta = base + 5 * nw; S.k = base + 6 * nw;
            S.reemission_cdf = base + 7 * nw;
            S.model = o->surface_model[s];
            S.transmissive = o->surface_transmissive[s];
            S.thickness = o->surface_thickness[s];
            S.wavelength_n = nw; S.wavelength_step = o->wavelength_step; S.wavelength_start = o->wavelength_start;
            const uint32_t na = o->dichroic_nangles ? o->dichroic_nangles[s] : 0;
            if (na) {
                DichroicProps &D = dichroic[s];
                D.angles = (float *)o->dichroic_angles + a0;
                for (uint32_t a = 0; a < na; a++) {
                    dich_ptrs[s * 2 + 0].push_back((float *)o->dichroic_reflect + (a0 + a) * nw);
                    dich_ptrs[s * 2 + 1].push_back((float *)o->dichroic_transmit + (a0 + a) * nw);
                }
                D.dichroic_reflect = dich_ptrs[s * 2 + 0].data();
                D.dichroic_transmit = dich_ptrs[s * 2 + 1].data();
                D.nangles = na;
                S.dichroic_props = &D;
            }
            a0 += na;
        }
        for (uint32_t s = 0; s < o->nsurfaces; s++)
            surface_ptrs.push_back((!o->surface_present || o->surface_present[s]) ? &surfaces[s] : nullptr);
        g.surfaces = surface_ptrs.data();
#ifdef REF_CPU_WITH_DETECTOR
        memset(&det, 0, sizeof(det));
        if (o->nchannels) {
            det.solid_id_to_channel_index = (int *)o->solid_id_to_channel_index;
            det.time_cdf_x = (float *)o->time_cdf_x; det.time_cdf_y = (float *)o->time_cdf_y;
            det.charge_cdf_x = (float *)o->charge_cdf_x; det.charge_cdf_y = (float *)o->charge_cdf_y;
            det.nchannels = (int)o->nchannels;
            det.time_cdf_len = (int)o->time_cdf_len; det.charge_cdf_len = (int)o->charge_cdf_len;
            det.charge_unit = o->charge_cdf_x[o->charge_cdf_len - 1] / 65536.0f;   // chroma/gpu/detector.py:39
        }
#endif
    }
};


}  // namespace
#endif
