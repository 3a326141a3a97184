// igmk.cu -- host side of the C ABI declared in include/igmk.h.
// Owns device memory, packs submitted structures into the flat batch layout of mstep.cuh,
// launches the kernels and copies results back.  No exceptions cross the ABI.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <cmath>
#include <string>
#include <vector>
#include <map>
#include <tuple>
#include <algorithm>
#include <chrono>
#include <thread>
#include <mutex>
#include <cuda_runtime.h>

#include "../../include/igmk.h"
#include "mstep.cuh"
#include "md2.cuh"
#include "md_big.cuh"
#include "md_cluster.cuh"
#include "astep.cuh"
#include "astep_shard.cuh"
#include "assemble.cuh"
#include "probe.cuh"

using namespace igmk;

static thread_local std::string g_err;

static int fail(const std::string& s) { g_err = s; return 1; }
static const int IGMK_NL_OVERFLOW = 100;   // internal: retry with a wider neighbour table

#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            char b_[512];                                                                     \
            snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return fail(b_);                                                                  \
        }                                                                                     \
    } while (0)

// copy on a stream and wait for it: unlike cudaMemcpy this never touches the legacy default stream, so batches
// running on other handles' streams are not serialised behind it
static cudaError_t copy_on(cudaStream_t st, void* dst, const void* src, size_t bytes, cudaMemcpyKind kind)
{
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, kind, st);
    return e != cudaSuccess ? e : cudaStreamSynchronize(st);
}

template <typename T>
static cudaError_t dev_alloc_copy(T** dptr, const std::vector<T>& h, cudaStream_t st)
{
    size_t bytes = std::max<size_t>(h.size(), 1) * sizeof(T);
    cudaError_t e = cudaMalloc((void**)dptr, bytes);
    if (e != cudaSuccess) return e;
    if (!h.empty()) e = copy_on(st, *dptr, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
    return e;
}

struct igmk_handle {
    int device = 0;
    int smem_optin = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;      // non-blocking: batches of different handles overlap on one device
    cudaStream_t stream_hi = nullptr;   // same, highest priority: the short CG launch that ends a batch overtakes the queued MD
                                        // CTAs of other batches, so a finished chunk leaves the device while the others still run
    // host staging
    std::vector<StructDesc> desc;
    std::vector<float4>   xyzr;         // current coordinates (results land here)
    std::vector<float4>   xyzr0;        // coordinates as submitted
    std::vector<uint8_t>  kind;
    std::vector<uint32_t> envmask;
    std::vector<int32_t>  bond_ptr;
    std::vector<uint32_t> bond_items;
    std::vector<float2>   btype_f;
    std::vector<double2>  btype_d;
    std::vector<int64_t>  struct_id;
    // bonds / envelope members in submit order, for the violation statistics (M10)
    std::vector<SubBond> sub_bonds; std::vector<int32_t> sub_bond_atom_off; std::vector<int64_t> sub_bond_start;
    std::vector<EnvMember> sub_env; std::vector<int64_t> sub_env_start;
    std::vector<double> viol_bond, viol_env; bool viol_valid = false;
    int max_atoms = 0;
    int nl_width = NL_WIDTH0;
    // thread-per-atom MD layout (md2.cuh), built at upload time
    int md2_K = 0, md2_nlw = MD2_NLW0;
    float skin_factor = 0.7f;   // k_md_big: single-level neighbour skin in bead radii
    float skin_cluster = 1.0f;  // k_md2c: its cluster-wide rebuild is dearer (0.6 / 0.7 / 0.8 / 1.0 -> 86 / 78 / 75 / 73 us per evaluation at 29 838 beads, tools/prof_c4.py 4 0.02)
    float skin_inner = 0.6f;    // k_md2: neighbour skin in bead radii (measured optimum once the rebuild got cheaper: 0.5 / 0.6 / 0.7 / 0.8 -> 73.8 / 72.4 / 73.6 / 75.2 ms, tools/prof_md.py 148 0.06)
    std::vector<Md2Desc> md2_desc;
    std::vector<uint16_t> md2_perm;
    std::vector<int32_t>  md2_ell_off;
    std::vector<int4>     md2_eblk;
    std::vector<uint32_t> md2_ell_items;
    std::string md2_error;
    int64_t md2_sym_bonds = 0, md2_bonds = 0;
    Md2Desc* d_md2_desc = nullptr; uint16_t* d_md2_perm = nullptr; int32_t* d_md2_ell_off = nullptr;
    uint32_t* d_md2_ell_items = nullptr; uint32_t* d_md2_ell_act = nullptr; uint16_t* d_md2_nl = nullptr;
    int4* d_md2_eblk = nullptr; float4* d_md2_gf = nullptr;
    // cluster path (md_cluster.cuh): CTAs per structure (0: not used) and its rebuild workspace
    int mdc_CL = 0;
    float4* d_mdc_xall = nullptr; float4* d_mdc_xsorted = nullptr; uint32_t* d_mdc_sorted = nullptr; uint32_t* d_mdc_cells = nullptr;
    uint16_t* d_mdc_rowlen = nullptr; float* d_mdc_bbox = nullptr;
    double* d_md2_dbg = nullptr;
    bool md_big = false; float* d_big_state = nullptr; double* d_big = nullptr;
    bool dirty = true;
    // batch assembled on the device (igmk_submit_population_dev, assemble.cuh): the host vectors above hold desc / struct_id /
    // xyzr only; the bond list stays on the device
    bool dev_built = false;
    std::vector<int32_t> dev_nb;                       // bonds per structure
    std::vector<double> dev_env_abck; std::vector<int32_t> dev_env_off, dev_env_mem;   // the template's envelopes (violation report)
    unsigned char* d_asm_blob = nullptr; int32_t* d_asm_cnt = nullptr;
    int32_t* d_asm_bl = nullptr; int32_t* d_asm_csr = nullptr; uint8_t* d_asm_dec = nullptr; uint16_t* d_asm_atom = nullptr;
    const double* d_asm_key_r0 = nullptr; const double* d_asm_key_k = nullptr; const int32_t* d_asm_key_style = nullptr;
    int64_t dev_bonds = 0;
    // device
    StructDesc* d_desc = nullptr; float4* d_xyzr = nullptr; float4* d_xyzr0 = nullptr; uint8_t* d_kind = nullptr; uint32_t* d_envmask = nullptr;
    int32_t* d_bond_ptr = nullptr; uint32_t* d_bond_items = nullptr; float2* d_btype_f = nullptr; double2* d_btype_d = nullptr;
    uint16_t* d_nl = nullptr; uint16_t* d_nl_cnt = nullptr; double* d_xd = nullptr; double* d_fd = nullptr; double* d_work = nullptr;
    igmk_result_t* d_res = nullptr; int32_t* d_flags = nullptr;
    // results
    std::vector<igmk_result_t> res;
    std::vector<int32_t> flags;
    std::vector<double> forces;
    bool have_forces = false;
    double md_ms = 0.0, cg_ms = 0.0;   // device time of the last igmk_run's two launches
    size_t h2d_bytes = 0;              // bytes the last upload() moved to the device

    void free_device()
    {
        cudaFree(d_desc); cudaFree(d_xyzr); cudaFree(d_xyzr0); cudaFree(d_kind); cudaFree(d_envmask); cudaFree(d_bond_ptr);
        cudaFree(d_bond_items); cudaFree(d_btype_f); cudaFree(d_btype_d); cudaFree(d_nl); cudaFree(d_nl_cnt);
        cudaFree(d_xd); cudaFree(d_fd); cudaFree(d_work); cudaFree(d_res); cudaFree(d_flags);
        cudaFree(d_big_state); cudaFree(d_big); d_big_state = nullptr; d_big = nullptr;
        cudaFree(d_md2_desc); cudaFree(d_md2_perm); cudaFree(d_md2_ell_off); cudaFree(d_md2_ell_items); cudaFree(d_md2_ell_act); cudaFree(d_md2_nl); d_md2_ell_act = nullptr;
        cudaFree(d_md2_eblk); cudaFree(d_md2_gf); cudaFree(d_md2_dbg);
        cudaFree(d_asm_blob); cudaFree(d_asm_cnt); cudaFree(d_asm_bl); cudaFree(d_asm_csr); cudaFree(d_asm_dec); cudaFree(d_asm_atom);
        d_asm_blob = nullptr; d_asm_cnt = nullptr; d_asm_bl = nullptr; d_asm_csr = nullptr; d_asm_dec = nullptr; d_asm_atom = nullptr;
        d_md2_eblk = nullptr; d_md2_gf = nullptr; d_md2_dbg = nullptr;
        cudaFree(d_mdc_xall); cudaFree(d_mdc_xsorted); cudaFree(d_mdc_sorted); cudaFree(d_mdc_cells); cudaFree(d_mdc_rowlen); cudaFree(d_mdc_bbox);
        d_mdc_xall = nullptr; d_mdc_xsorted = nullptr; d_mdc_sorted = nullptr; d_mdc_cells = nullptr; d_mdc_rowlen = nullptr; d_mdc_bbox = nullptr;
        d_md2_desc = nullptr; d_md2_perm = nullptr; d_md2_ell_off = nullptr; d_md2_ell_items = nullptr; d_md2_nl = nullptr;
        d_desc = nullptr; d_xyzr = nullptr; d_xyzr0 = nullptr; d_kind = nullptr; d_envmask = nullptr; d_bond_ptr = nullptr;
        d_bond_items = nullptr; d_btype_f = nullptr; d_btype_d = nullptr; d_nl = nullptr; d_nl_cnt = nullptr;
        d_xd = nullptr; d_fd = nullptr; d_work = nullptr; d_res = nullptr; d_flags = nullptr;
    }
};

extern "C" const char* igmk_last_error(void) { return g_err.c_str(); }
extern "C" int igmk_version(void) { return 100; }
extern "C" int igmk_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int igmk_create(int device, igmk_handle_t** out)
{
    if (!out) return fail("igmk_create: out is NULL");
    if (device == IGMK_STAGING_ONLY) {          // host staging without a GPU: submit / slot_sizes / layout_stats only
        igmk_handle* h = new igmk_handle();
        h->device = IGMK_STAGING_ONLY;
        h->smem_optin = 227 * 1024;             // the B200 opt-in limit, for the same size checks as on the device
        *out = h;
        return 0;
    }
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail("igmk_create: no such CUDA device");
    CU(cudaSetDevice(device));
    igmk_handle* h = new igmk_handle();
    h->device = device;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    h->smem_optin = (int)prop.sharedMemPerBlockOptin;
    h->sm_count = prop.multiProcessorCount;
    if (const char* e = getenv("IGMK_NL_WIDTH")) {          // initial neighbour-table width (tuning / profiling aid)
        const int w = atoi(e);
        if (w >= 8 && w <= 65536) { h->md2_nlw = (w + 7) / 8 * 8; h->nl_width = h->md2_nlw; }
    }
    if (const char* e = getenv("IGMK_SKIN")) {              // neighbour skin in units of the bead radius (results do not depend on it)
        const float f = (float)atof(e);
        if (f >= 0.05f && f <= 4.0f) { h->skin_factor = f; h->skin_inner = f; h->skin_cluster = f; }
    }

    CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    int prio_lo = 0, prio_hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    CU(cudaStreamCreateWithPriority(&h->stream_hi, cudaStreamNonBlocking, prio_hi));
    *out = h;
    return 0;
}

extern "C" int igmk_destroy(igmk_handle_t* h)
{
    if (!h) return 0;
    if (h->device < 0) { delete h; return 0; }
    cudaSetDevice(h->device);
    h->free_device();
    if (h->stream) cudaStreamDestroy(h->stream);
    if (h->stream_hi) cudaStreamDestroy(h->stream_hi);
    delete h;
    return 0;
}

extern "C" int igmk_clear(igmk_handle_t* h)
{
    if (!h) return fail("igmk_clear: NULL handle");
    if (h->device >= 0) {
        cudaSetDevice(h->device);
        h->free_device();
    }
    h->desc.clear(); h->xyzr.clear(); h->xyzr0.clear(); h->kind.clear(); h->envmask.clear(); h->bond_ptr.clear();
    h->bond_items.clear(); h->btype_f.clear(); h->btype_d.clear(); h->struct_id.clear();
    h->sub_bonds.clear(); h->sub_bond_atom_off.clear(); h->sub_bond_start.clear(); h->sub_env.clear(); h->sub_env_start.clear();
    h->viol_valid = false;
    h->res.clear(); h->flags.clear(); h->forces.clear();
    h->max_atoms = 0; h->dirty = true; h->have_forces = false;
    h->dev_built = false; h->dev_nb.clear(); h->dev_bonds = 0;
    return 0;
}

extern "C" int igmk_submit(igmk_handle_t* h, const igmk_system_t* s, int32_t* slot_out)
{
    if (!h || !s) return fail("igmk_submit: NULL argument");
    if (h->dev_built) return fail("igmk_submit: this batch was assembled on the device; igmk_clear it first");
    char msg[256];
    const int n = s->n_atoms;
    if (n <= 0 || n > 65534) {
        snprintf(msg, sizeof msg, "igmk_submit: structure %lld has %d atoms (supported: 1..65534)", (long long)s->struct_id, n);
        return fail(msg);
    }
    if (s->n_env < 0 || s->n_env > IGMK_MAX_ENVELOPES) return fail("igmk_submit: too many envelopes");
    const size_t need = std::max(ctx_smem_bytes<float>(n, true), ctx_smem_bytes<double>(n, false));
    if ((long long)need > (long long)h->smem_optin) {
        snprintf(msg, sizeof msg, "igmk_submit: structure %lld with %d atoms needs %zu B of shared memory per CTA, "
                 "device offers %d B (single-CTA path; larger structures need the cluster path)",
                 (long long)s->struct_id, n, need, h->smem_optin);
        return fail(msg);
    }
    StructDesc d;
    memset(&d, 0, sizeof d);
    d.atom_off = (int32_t)h->xyzr.size();
    d.n_atoms = n;
    d.bond_ptr_off = (int32_t)h->bond_ptr.size();
    d.bond_item_off = (int64_t)h->bond_items.size();
    d.btype_off = (int32_t)h->btype_f.size();
    d.n_env = s->n_env;
    d.evfactor = s->evfactor;
    d.seed = s->seed;
    float maxrad = 0.f;
    int nonfixed = 0;
    for (int i = 0; i < n; i++) {
        if (s->kind[i] > 2) return fail("igmk_submit: unknown particle kind");
        h->xyzr.push_back(make_float4(s->xyz[3 * i], s->xyz[3 * i + 1], s->xyz[3 * i + 2], s->radius[i]));
        h->xyzr0.push_back(h->xyzr.back());
        h->kind.push_back(s->kind[i]);
        if (s->kind[i] == IGMK_KIND_BEAD) maxrad = std::max(maxrad, s->radius[i]);
        if (s->kind[i] != IGMK_KIND_STATIC) nonfixed++;
    }
    d.maxrad = maxrad;
    d.n_nonfixed = nonfixed;
    // envelopes -> per-atom membership mask
    std::vector<uint32_t> mask(n, 0u);
    for (int e = 0; e < s->n_env; e++) {
        for (int k = 0; k < 4; k++) d.env[e][k] = s->env_abck[4 * e + k];
        for (int q = s->env_off[e]; q < s->env_off[e + 1]; q++) {
            int i = s->env_members[q];
            if (i < 0 || i >= n) return fail("igmk_submit: envelope member out of range");
            mask[i] |= 1u << e;
        }
    }
    h->envmask.insert(h->envmask.end(), mask.begin(), mask.end());
    // bonds -> directional CSR by atom, bond types de-duplicated like LammpsModel.add_bond
    // (lammps_model.py:314-329) so that the type table stays cache resident
    std::map<std::tuple<double, double, int>, int> types;
    std::vector<int32_t> cnt(n + 1, 0);
    for (int b = 0; b < s->n_bonds; b++) {
        int i = s->bond_i[b], j = s->bond_j[b];
        if (i < 0 || i >= n || j < 0 || j >= n) return fail("igmk_submit: bond index out of range");
        if (s->bond_style[b] > 1) return fail("igmk_submit: unknown bond style");
        if (!(s->bond_k[b] >= 0.0)) return fail("igmk_submit: bond spring constants must be >= 0");
        if (!(s->bond_r0[b] >= 0.0)) {      // the kernels compare squared distances; a negative bound (FISH target < tolerance, fish.py:193) has no meaning for them
            snprintf(msg, sizeof msg, "igmk_submit: bond %d of structure %lld has a negative bound distance %g", b, (long long)s->struct_id, s->bond_r0[b]);
            return fail(msg);
        }
        cnt[i + 1]++; cnt[j + 1]++;
    }
    for (int i = 0; i < n; i++) cnt[i + 1] += cnt[i];
    std::vector<uint32_t> items((size_t)cnt[n]);
    std::vector<int32_t> cur(cnt.begin(), cnt.end() - 1);
    // provisional type ids in first-seen order, then renumbered by use count so that the MD kernel
    // can keep the two most frequent types in registers
    std::vector<int32_t> btype_of((size_t)s->n_bonds);
    std::vector<std::tuple<double, double, int>> tkeys;
    std::vector<int64_t> tuse;
    int last_t = -1;
    for (int b = 0; b < s->n_bonds; b++) {
        if (last_t >= 0 && s->bond_r0[b] == s->bond_r0[b - 1] && s->bond_k[b] == s->bond_k[b - 1] && s->bond_style[b] == s->bond_style[b - 1]) {
            tuse[last_t]++; btype_of[b] = last_t;       // runs of equal restraints (Hi-C, polymer) skip the map
            continue;
        }
        auto key = std::make_tuple(s->bond_r0[b], s->bond_k[b], (int)s->bond_style[b]);
        auto it = types.find(key);
        int t;
        if (it == types.end()) {
            t = (int)types.size();
            if (t > 65535) return fail("igmk_submit: more than 65536 distinct bond types in one structure");
            types[key] = t;
            tkeys.push_back(key); tuse.push_back(0);
        } else t = it->second;
        tuse[t]++;
        btype_of[b] = t;
        last_t = t;
    }
    std::vector<int32_t> torder(tkeys.size()), tnew(tkeys.size());
    for (size_t t = 0; t < torder.size(); t++) torder[t] = (int32_t)t;
    std::stable_sort(torder.begin(), torder.end(), [&](int32_t a, int32_t b) { return tuse[a] > tuse[b]; });
    for (size_t r = 0; r < torder.size(); r++) {
        tnew[torder[r]] = (int32_t)r;
        const double r0 = std::get<0>(tkeys[torder[r]]), k = std::get<1>(tkeys[torder[r]]);
        const double ks = std::get<2>(tkeys[torder[r]]) == IGMK_BOND_LOWER ? -k : k;
        h->btype_f.push_back(make_float2((float)r0, (float)ks));
        h->btype_d.push_back(make_double2(r0, ks));
    }
    for (size_t r = torder.size(); r < 2; r++) {           // the kernel reads entries 0 and 1 unconditionally
        h->btype_f.push_back(make_float2(0.f, 0.f));
        h->btype_d.push_back(make_double2(0.0, 0.0));
    }
    for (int b = 0; b < s->n_bonds; b++) {
        const uint32_t t = (uint32_t)tnew[btype_of[b]];
        const int i = s->bond_i[b], j = s->bond_j[b];
        items[cur[i]++] = (uint32_t)j | (t << 16);
        items[cur[j]++] = (uint32_t)i | (t << 16);
    }
    h->bond_ptr.insert(h->bond_ptr.end(), cnt.begin(), cnt.end());
    h->bond_items.insert(h->bond_items.end(), items.begin(), items.end());
    h->desc.push_back(d);
    h->struct_id.push_back(s->struct_id);
    h->sub_bond_start.push_back((int64_t)h->sub_bonds.size());
    for (int b = 0; b < s->n_bonds; b++) {
        SubBond q;
        q.i = s->bond_i[b]; q.j = s->bond_j[b]; q.r0 = s->bond_r0[b]; q.k = s->bond_k[b]; q.style = s->bond_style[b];
        q.f32 = ((double)(float)q.r0 == q.r0) ? 1 : 0;
        h->sub_bonds.push_back(q);
        h->sub_bond_atom_off.push_back(d.atom_off);
    }
    h->sub_env_start.push_back((int64_t)h->sub_env.size());
    for (int e = 0; e < s->n_env; e++)
        for (int q = s->env_off[e]; q < s->env_off[e + 1]; q++) {
            EnvMember m;
            m.atom = d.atom_off + s->env_members[q]; m.pad = 0;
            m.a = s->env_abck[4 * e]; m.b = s->env_abck[4 * e + 1]; m.c = s->env_abck[4 * e + 2]; m.k = s->env_abck[4 * e + 3];
            h->sub_env.push_back(m);
        }
    h->viol_valid = false;
    h->max_atoms = std::max(h->max_atoms, n);
    h->dirty = true;
    if (slot_out) *slot_out = (int32_t)h->desc.size() - 1;
    return 0;
}

// ---- measured rates for the MD kernel's roofline (probe.cuh) ----------------------------------
extern "C" int igmk_peak_probe(int device, double* fp32_tflops, double* fp32x2_tflops, double* smem_tbs)
{
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail("igmk_peak_probe: no such CUDA device");
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    const int grid = prop.multiProcessorCount;          // one 1024-thread CTA per SM, like k_md2
    float* d_out = nullptr;
    CU(cudaMalloc((void**)&d_out, 16));
    cudaStream_t st;
    CU(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    const size_t lds_smem = 4 * PROBE_THREADS * sizeof(float4);
    CU(cudaFuncSetAttribute(k_probe_lds, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lds_smem));
    const int it_f = 1 << 17, it_l = 1 << 15;
    double best[3] = {0, 0, 0};
    for (int rep = 0; rep < 6; rep++)                    // rep 0 warms up; best of the rest
        for (int which = 0; which < 3; which++) {
            CU(cudaEventRecord(e0, st));
            if (which == 0) k_probe_ffma<<<grid, PROBE_THREADS, 0, st>>>(it_f, 1.000001f, 1e-7f, d_out);
            else if (which == 1) k_probe_ffma2<<<grid, PROBE_THREADS, 0, st>>>(it_f, 1.000001f, 1e-7f, d_out);
            else k_probe_lds<<<grid, PROBE_THREADS, lds_smem, st>>>(it_l, d_out);
            CU(cudaGetLastError());
            CU(cudaEventRecord(e1, st));
            CU(cudaEventSynchronize(e1));
            float ms = 0.f;
            CU(cudaEventElapsedTime(&ms, e0, e1));
            if (rep == 0 || ms <= 0.f) continue;
            const double per_s = 1e3 / ms;
            const double v = which < 2 ? (double)grid * PROBE_THREADS * PROBE_CHAINS * 2.0 * it_f * per_s / 1e12
                                       : (double)grid * PROBE_THREADS * 4.0 * 16.0 * it_l * per_s / 1e12;
            best[which] = std::max(best[which], v);
        }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaStreamDestroy(st); cudaFree(d_out);
    if (fp32_tflops) *fp32_tflops = best[0];
    if (fp32x2_tflops) *fp32x2_tflops = best[1];
    if (smem_tbs) *smem_tbs = best[2];
    return 0;
}

// ---- whole population from a template + the assignment bit matrix ---------------------------
// Append the staged structures of `parts` (in this order) to dst: the destination vectors grow once, then
// every part copies its slice on its own thread, rebasing its offsets.
template <typename T>
static size_t grow(std::vector<T>& v, size_t n) { const size_t at = v.size(); v.resize(at + n); return at; }

static void append_parts(igmk_handle* dst, const std::vector<igmk_handle*>& parts)
{
    struct Base { size_t desc, atoms, bptr, bitems, btype, sbond, senv; };
    std::vector<Base> base(parts.size());
    Base tot = {0, 0, 0, 0, 0, 0, 0};
    for (size_t t = 0; t < parts.size(); t++) {
        const igmk_handle* p = parts[t];
        base[t] = tot;
        tot.desc += p->desc.size(); tot.atoms += p->xyzr.size(); tot.bptr += p->bond_ptr.size(); tot.bitems += p->bond_items.size();
        tot.btype += p->btype_f.size(); tot.sbond += p->sub_bonds.size(); tot.senv += p->sub_env.size();
        dst->max_atoms = std::max(dst->max_atoms, p->max_atoms);
    }
    const Base at = {grow(dst->desc, tot.desc), grow(dst->xyzr, tot.atoms), grow(dst->bond_ptr, tot.bptr), grow(dst->bond_items, tot.bitems),
                     grow(dst->btype_f, tot.btype), grow(dst->sub_bonds, tot.sbond), grow(dst->sub_env, tot.senv)};
    grow(dst->xyzr0, tot.atoms); grow(dst->kind, tot.atoms); grow(dst->envmask, tot.atoms); grow(dst->btype_d, tot.btype);
    grow(dst->struct_id, tot.desc); grow(dst->sub_bond_start, tot.desc); grow(dst->sub_env_start, tot.desc);
    grow(dst->sub_bond_atom_off, tot.sbond);
    std::vector<std::thread> pool;
    for (size_t t = 0; t < parts.size(); t++)
        pool.emplace_back([&, t]() {
            const igmk_handle* p = parts[t];
            const Base o = {at.desc + base[t].desc, at.atoms + base[t].atoms, at.bptr + base[t].bptr, at.bitems + base[t].bitems,
                            at.btype + base[t].btype, at.sbond + base[t].sbond, at.senv + base[t].senv};
            for (size_t k = 0; k < p->desc.size(); k++) {
                StructDesc d = p->desc[k];
                d.atom_off += (int32_t)o.atoms; d.bond_ptr_off += (int32_t)o.bptr; d.bond_item_off += (int64_t)o.bitems; d.btype_off += (int32_t)o.btype;
                dst->desc[o.desc + k] = d;
                dst->struct_id[o.desc + k] = p->struct_id[k];
                dst->sub_bond_start[o.desc + k] = p->sub_bond_start[k] + (int64_t)o.sbond;
                dst->sub_env_start[o.desc + k] = p->sub_env_start[k] + (int64_t)o.senv;
            }
            std::copy(p->xyzr.begin(), p->xyzr.end(), dst->xyzr.begin() + o.atoms);
            std::copy(p->xyzr0.begin(), p->xyzr0.end(), dst->xyzr0.begin() + o.atoms);
            std::copy(p->kind.begin(), p->kind.end(), dst->kind.begin() + o.atoms);
            std::copy(p->envmask.begin(), p->envmask.end(), dst->envmask.begin() + o.atoms);
            std::copy(p->bond_ptr.begin(), p->bond_ptr.end(), dst->bond_ptr.begin() + o.bptr);
            std::copy(p->bond_items.begin(), p->bond_items.end(), dst->bond_items.begin() + o.bitems);
            std::copy(p->btype_f.begin(), p->btype_f.end(), dst->btype_f.begin() + o.btype);
            std::copy(p->btype_d.begin(), p->btype_d.end(), dst->btype_d.begin() + o.btype);
            std::copy(p->sub_bonds.begin(), p->sub_bonds.end(), dst->sub_bonds.begin() + o.sbond);
            for (size_t k = 0; k < p->sub_bond_atom_off.size(); k++) dst->sub_bond_atom_off[o.sbond + k] = p->sub_bond_atom_off[k] + (int32_t)o.atoms;
            for (size_t k = 0; k < p->sub_env.size(); k++) { EnvMember m = p->sub_env[k]; m.atom += (int32_t)o.atoms; dst->sub_env[o.senv + k] = m; }
        });
    for (auto& th : pool) th.join();
    dst->dirty = true; dst->viol_valid = false;
}

extern "C" int igmk_submit_population(igmk_handle_t* h, const igmk_system_t* tmpl, int32_t n_bead, int32_t n_struct_total,
                                      int32_t s_begin, int32_t n_struct, const float* crd,
                                      int64_t n_rows, const int32_t* row, const int32_t* col, const double* row_r0, double row_k,
                                      const uint32_t* bits, int32_t bits_words, const int64_t* seeds, int32_t* first_slot_out)
{
    return igmk_submit_population_ex(h, tmpl, n_bead, n_struct_total, s_begin, n_struct, crd, n_rows, row, col, row_r0, row_k,
                                     bits, bits_words, seeds, nullptr, first_slot_out);
}

extern "C" int igmk_submit_population_ex(igmk_handle_t* h, const igmk_system_t* tmpl, int32_t n_bead, int32_t n_struct_total,
                                         int32_t s_begin, int32_t n_struct, const float* crd,
                                         int64_t n_rows, const int32_t* row, const int32_t* col, const double* row_r0, double row_k,
                                         const uint32_t* bits, int32_t bits_words, const int64_t* seeds,
                                         const igmk_pop_extras_t* ex, int32_t* first_slot_out)
{
    if (!h || !tmpl || !crd) return fail("igmk_submit_population: NULL argument");
    if (ex) {
        if (ex->n_env < 0 || tmpl->n_env + ex->n_env > IGMK_MAX_ENVELOPES) return fail("igmk_submit_population: too many envelopes");
        if (ex->n_env > 0 && (!ex->env_abck || !ex->env_row_off || (ex->env_row_off[ex->n_env] > 0 && (!ex->env_bead || !ex->env_bits))))
            return fail("igmk_submit_population: NULL envelope arguments in extras");
        if (ex->n_env > 0 && ex->env_row_off[ex->n_env] > 0 && ex->env_bits_words * 32 < n_struct)
            return fail("igmk_submit_population: envelope bit matrix narrower than the structure range");
        if (ex->bond_off && ex->bond_off[n_struct] > ex->bond_off[0] &&
            (!ex->bond_i || !ex->bond_j || !ex->bond_k || !ex->bond_r0 || !ex->bond_style))
            return fail("igmk_submit_population: NULL bond arguments in extras");
        if (ex->atom_off && ex->atom_off[n_struct] > ex->atom_off[0] && (!ex->atom_xyz || !ex->atom_radius || !ex->atom_kind))
            return fail("igmk_submit_population: NULL atom arguments in extras");
    }
    if (n_bead < 1 || n_bead > tmpl->n_atoms) return fail("igmk_submit_population: n_bead must be 1..tmpl->n_atoms");
    if (n_struct < 0 || s_begin < 0 || s_begin + n_struct > n_struct_total) return fail("igmk_submit_population: bad structure range");
    if (n_rows > 0 && (!row || !col || !row_r0 || !bits)) return fail("igmk_submit_population: NULL row arguments");
    if (n_rows > 0 && bits_words * 32 < n_struct) return fail("igmk_submit_population: bit matrix narrower than the structure range");
    if (first_slot_out) *first_slot_out = (int32_t)h->desc.size();
    if (n_struct == 0) return 0;
    const bool trace_sp = getenv("IGMK_TRACE") != nullptr;
    auto tnow = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double tsp0 = tnow();
    // Rows of every structure, in row order (bit s of word s/32 = local structure s).  The row range is cut into
    // one slice per thread; each thread validates its rows and lists them per structure; the slices are then
    // concatenated in slice order, which is row order.
    const int nscan = (int)std::max<int64_t>(1, std::min<int64_t>(std::thread::hardware_concurrency(), n_rows / 65536 + 1));
    std::vector<std::vector<std::vector<int32_t>>> found((size_t)nscan, std::vector<std::vector<int32_t>>((size_t)n_struct));
    std::vector<int> bad((size_t)nscan, 0);
    {
        std::vector<std::thread> scan;
        for (int t = 0; t < nscan; t++)
            scan.emplace_back([&, t]() {
                const int64_t r0 = n_rows * t / nscan, r1 = n_rows * (t + 1) / nscan;
                auto& mine_t = found[t];
                for (int64_t r = r0; r < r1; r++) {
                    if (row[r] < 0 || row[r] >= n_bead || col[r] < 0 || col[r] >= n_bead) { bad[t] = 1; return; }
                    const uint32_t* w = bits + (size_t)r * bits_words;
                    for (int k = 0; k * 32 < n_struct; k++) {
                        uint32_t m = w[k];
                        while (m) {
                            const int b = __builtin_ctz(m); m &= m - 1;
                            const int sl = k * 32 + b;
                            if (sl < n_struct) mine_t[sl].push_back((int32_t)r);
                        }
                    }
                }
            });
        for (auto& th : scan) th.join();
    }
    for (int t = 0; t < nscan; t++) if (bad[t]) return fail("igmk_submit_population: row bead index out of range");
    const double tsp1 = tnow();
    const int nthr = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), (unsigned)n_struct));
    std::vector<igmk_handle*> part((size_t)nthr, nullptr);
    std::vector<std::string> errs((size_t)nthr);
    std::vector<std::thread> pool;
    for (int t = 0; t < nthr; t++) {
        part[t] = new igmk_handle();
        part[t]->smem_optin = h->smem_optin;
        pool.emplace_back([&, t]() {
            const int lo = (int)((int64_t)n_struct * t / nthr), hi = (int)((int64_t)n_struct * (t + 1) / nthr);
            const int na = tmpl->n_atoms;
            std::vector<float> xyz((size_t)na * 3);
            std::vector<int32_t> bi, bj; std::vector<double> bk, br; std::vector<uint8_t> bs;
            std::vector<double> eabck; std::vector<int32_t> eoff, emem;
            {                                             // this thread's share of the staging vectors, allocated once
                size_t nb = 0;
                for (int sl = lo; sl < hi; sl++) { nb += (size_t)tmpl->n_bonds; for (int q = 0; q < nscan; q++) nb += found[q][sl].size(); }
                igmk_handle* ph = part[t];
                const size_t nat = (size_t)(hi - lo) * na;
                ph->xyzr.reserve(nat); ph->xyzr0.reserve(nat); ph->kind.reserve(nat); ph->envmask.reserve(nat);
                ph->bond_ptr.reserve((size_t)(hi - lo) * (na + 1)); ph->bond_items.reserve(2 * nb);
                ph->sub_bonds.reserve(nb); ph->sub_bond_atom_off.reserve(nb);
            }
            std::vector<float> xrad; std::vector<uint8_t> xkind;
            for (int sl = lo; sl < hi; sl++) {
                const int sg = s_begin + sl;
                const int nx = (ex && ex->atom_off) ? (int)(ex->atom_off[sl + 1] - ex->atom_off[sl]) : 0;     // this structure's own particles
                xyz.resize((size_t)(na + nx) * 3);
                for (int a = 0; a < n_bead; a++) {
                    const float* src = crd + ((size_t)a * n_struct_total + sg) * 3;
                    xyz[3 * a] = src[0]; xyz[3 * a + 1] = src[1]; xyz[3 * a + 2] = src[2];
                }
                for (int a = n_bead; a < na; a++) { xyz[3 * a] = tmpl->xyz[3 * a]; xyz[3 * a + 1] = tmpl->xyz[3 * a + 1]; xyz[3 * a + 2] = tmpl->xyz[3 * a + 2]; }
                if (nx > 0) {                          // e.g. the centroid particles of the SPRITE clusters assigned to this structure (sprite.py:57-58)
                    const int64_t o = ex->atom_off[sl];
                    memcpy(xyz.data() + (size_t)na * 3, ex->atom_xyz + o * 3, (size_t)nx * 3 * sizeof(float));
                    xrad.assign(tmpl->radius, tmpl->radius + na); xrad.insert(xrad.end(), ex->atom_radius + o, ex->atom_radius + o + nx);
                    xkind.assign(tmpl->kind, tmpl->kind + na); xkind.insert(xkind.end(), ex->atom_kind + o, ex->atom_kind + o + nx);
                }
                bi.assign(tmpl->bond_i, tmpl->bond_i + tmpl->n_bonds); bj.assign(tmpl->bond_j, tmpl->bond_j + tmpl->n_bonds);
                bk.assign(tmpl->bond_k, tmpl->bond_k + tmpl->n_bonds); br.assign(tmpl->bond_r0, tmpl->bond_r0 + tmpl->n_bonds);
                bs.assign(tmpl->bond_style, tmpl->bond_style + tmpl->n_bonds);
                for (int q = 0; q < nscan; q++)
                    for (int32_t r : found[q][sl]) {   // HarmonicUpperBound((row, col), r0, k), inter_hic.py:55-62
                        bi.push_back(row[r]); bj.push_back(col[r]); bk.push_back(row_k); br.push_back(row_r0[r]); bs.push_back(IGMK_BOND_UPPER);
                    }
                igmk_system_t sys = *tmpl;
                if (ex && ex->bond_off)                // this structure's own restraints (FISH, fish.py:93-274), after the Hi-C ones
                    for (int64_t q = ex->bond_off[sl]; q < ex->bond_off[sl + 1]; q++) {
                        bi.push_back(ex->bond_i[q]); bj.push_back(ex->bond_j[q]); bk.push_back(ex->bond_k[q]);
                        br.push_back(ex->bond_r0[q]); bs.push_back(ex->bond_style[q]);
                    }
                if (ex && ex->n_env > 0) {             // envelope groups whose members differ by structure (DamID, damid.py:116-142)
                    eabck.assign(tmpl->env_abck, tmpl->env_abck + 4 * (size_t)tmpl->n_env);
                    eabck.insert(eabck.end(), ex->env_abck, ex->env_abck + 4 * (size_t)ex->n_env);
                    eoff.assign(tmpl->env_off, tmpl->env_off + tmpl->n_env + 1);
                    emem.assign(tmpl->env_members, tmpl->env_members + tmpl->env_off[tmpl->n_env]);
                    for (int e = 0; e < ex->n_env; e++) {
                        for (int64_t r = ex->env_row_off[e]; r < ex->env_row_off[e + 1]; r++)
                            if ((ex->env_bits[(size_t)r * ex->env_bits_words + (sl >> 5)] >> (sl & 31)) & 1u) emem.push_back(ex->env_bead[r]);
                        eoff.push_back((int32_t)emem.size());
                    }
                    sys.n_env = tmpl->n_env + ex->n_env;
                    sys.env_abck = eabck.data(); sys.env_off = eoff.data(); sys.env_members = emem.data();
                }
                sys.xyz = xyz.data();
                if (nx > 0) { sys.n_atoms = na + nx; sys.radius = xrad.data(); sys.kind = xkind.data(); }
                sys.n_bonds = (int32_t)bi.size();
                sys.bond_i = bi.data(); sys.bond_j = bj.data(); sys.bond_k = bk.data(); sys.bond_r0 = br.data(); sys.bond_style = bs.data();
                sys.struct_id = tmpl->struct_id + sg;
                // no seed list: the script's formula (lammps.py:159-165) with tmpl->seed standing for seed * step_no, so that
                // every structure draws its own velocities (structure id 0 gets seed 1, as in the reference)
                sys.seed = seeds ? seeds[sl] : (int64_t)(((__int128)std::max<int64_t>(tmpl->seed, 1) * sys.struct_id) % 9190037) + 1;
                if (igmk_submit(part[t], &sys, nullptr)) { errs[t] = igmk_last_error(); return; }
            }
        });
    }
    for (auto& th : pool) th.join();
    const double tsp2 = tnow();
    int rc = 0;
    for (int t = 0; t < nthr; t++) if (!errs[t].empty() && !rc) rc = fail(errs[t]);
    if (!rc) append_parts(h, part);
    for (int t = 0; t < nthr; t++) delete part[t];
    if (trace_sp)
        fprintf(stderr, "[igmk trace] submit_population of %d: bit scan %.1f ms, per-structure submits %.1f ms, merge %.1f ms\n",
                n_struct, tsp1 - tsp0, tsp2 - tsp1, tnow() - tsp2);
    return rc;
}

// slots per thread the MD kernel is instantiated for
static int md2_pick_K(int max_atoms)
{
    const int need = (max_atoms + MD2S_T - 1) / MD2S_T;
    const int avail[] = {1, 2, 3, 4, 5};     // 5 slots x 768 threads = 3840 atoms is what one SM's 227 KB of shared memory holds (md2_smem_bytes)
    for (int k : avail) if (k >= need) return k;
    return 0;                                  // larger structures: a thread-block cluster per structure or the global-memory kernel
}

// Atom ordering + sliced-ELL bond items for the thread-per-atom MD kernel: atoms sorted by bond
// count, 32 at a time into warp-slots laid out boustrophedon over the warps so that every warp
// gets a similar amount of bond work and the atoms of one warp-slot have similar row lengths.
static void md2_build(igmk_handle* h)
{
    const int K = h->md2_K, np = K * MD2_T, NW = MD2_NW, G = K * NW;
    const size_t ns = h->desc.size();
    h->md2_desc.assign(ns, Md2Desc());
    h->md2_perm.assign(ns * (size_t)np, MD2_NONE);
    h->md2_ell_off.assign(ns * (size_t)(G + 1), 0);
    std::vector<std::vector<uint32_t>> items(ns);            // per structure, concatenated below
    auto build_one = [&](size_t s) {
        std::vector<int> order, pos;
        const StructDesc& d = h->desc[s];
        const int n = d.n_atoms;
        const int32_t* bp = h->bond_ptr.data() + d.bond_ptr_off;
        const uint32_t* bi = h->bond_items.data() + d.bond_item_off;
        order.resize(n); pos.assign(n, -1);
        for (int a = 0; a < n; a++) order[a] = a;
        // Large structures (state in global memory) keep the sequence order: consecutive beads are neighbours in
        // space, so the position gathers of a warp fall into fewer L2 sectors; small ones (state in shared memory)
        // are ordered by bond count, which minimises the padding of the sliced-ELL blocks.
        if (!h->md_big)
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return bp[a + 1] - bp[a] > bp[b + 1] - bp[b]; });
        uint16_t* perm = h->md2_perm.data() + s * (size_t)np;
        auto slot_of_group = [&](int g) { const int k = g / NW, r = g % NW; return k * NW + ((k & 1) ? NW - 1 - r : r); };
        for (int g = 0; g < G; g++) {
            const int gi = slot_of_group(g), k = gi / NW, w = gi % NW;
            for (int l = 0; l < 32; l++) {
                const int q = g * 32 + l;
                if (q >= n) break;
                const int p = k * MD2_T + w * 32 + l;
                perm[p] = (uint16_t)order[q]; pos[order[q]] = p;
            }
        }
        Md2Desc& m = h->md2_desc[s];
        m.perm_off = (int32_t)(s * (size_t)np);
        m.ell_desc_off = (int32_t)(s * (size_t)(G + 1));
        m.K = K;
        int32_t* eoff = h->md2_ell_off.data() + m.ell_desc_off;
        std::vector<uint32_t>& out = items[s];
        for (int gi = 0; gi < G; gi++) {           // blocks in (slot, warp) order = the order the kernel indexes them
            const int k = gi / NW, w = gi % NW;
            int cols = 0;
            for (int l = 0; l < 32; l++) {
                const int a = perm[k * MD2_T + w * 32 + l];
                if (a != MD2_NONE) cols = std::max(cols, bp[a + 1] - bp[a]);
            }
            cols = (cols + 3) / 4 * 4;                 // the kernel reads 4 columns at a time, unguarded
            eoff[gi] = (int32_t)out.size();
            out.resize(out.size() + (size_t)cols * 32, 0u);
            uint32_t* blk = out.data() + eoff[gi];
            // layout [batch of 4 columns][lane][4]: one 16-byte load per lane and batch
            auto at = [&](int c, int l) -> uint32_t& { return blk[((c >> 2) * 32 + l) * 4 + (c & 3)]; };
            for (int l = 0; l < 32; l++)               // padding = bond of type 0 to oneself: r = 0, no force
                for (int c = 0; c < cols; c++) at(c, l) = (uint32_t)(k * MD2_T + w * 32 + l);
            for (int l = 0; l < 32; l++) {
                const int a = perm[k * MD2_T + w * 32 + l];
                if (a == MD2_NONE) continue;
                for (int c = 0; c < bp[a + 1] - bp[a]; c++) {
                    const uint32_t it = bi[bp[a] + c];
                    at(c, l) = (uint32_t)pos[it & 0xffffu] | (it & 0xffff0000u);
                }
            }
        }
        eoff[G] = (int32_t)out.size();
    };
    const int nthr = (int)std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), ns));
    std::vector<std::thread> pool;
    for (int t = 0; t < nthr; t++)
        pool.emplace_back([&, t]() { for (size_t s = (size_t)t; s < ns; s += (size_t)nthr) build_one(s); });
    for (auto& th : pool) th.join();
    size_t total = 0;
    for (size_t s = 0; s < ns; s++) { h->md2_desc[s].ell_item_off = (int64_t)total; total += items[s].size(); }
    h->md2_ell_items.resize(total);
    pool.clear();
    for (int t = 0; t < nthr; t++)
        pool.emplace_back([&, t]() {
            for (size_t s = (size_t)t; s < ns; s += (size_t)nthr)
                if (!items[s].empty())
                    memcpy(h->md2_ell_items.data() + h->md2_desc[s].ell_item_off, items[s].data(), items[s].size() * sizeof(uint32_t));
        });
    for (auto& th : pool) th.join();
}

// Layout for k_md2 (md2.cuh): every bond is ORIENTED -- one end owns it, the other receives it through an inbox slot --
// and the owned items of an atom are split into block A (the most frequent bond type, when that is an upper bound) and
// block B (everything else, and bonds whose two ends both have full inboxes: listed at both ends, SYM).  Atoms are
// ordered by the number of items they own, 32 at a time into warp-slots laid out boustrophedon over the warps.
// CL > 1: the layout of k_md2c (md_cluster.cuh) -- the atoms are cut into CL contiguous blocks in index order (genome
// order: a chromosome territory is compact in space), block r goes to CTA r of the cluster, a position is r * np + offset
// and the item fields are wider.
static void md2_build_owner(igmk_handle* h, int CL = 1)
{
    const int T = CL > 1 ? MD2_T : MD2S_T;                           // threads per CTA: k_md2c / k_md2
    const int K = h->md2_K, npc = K * T, np = npc * CL, NW = T / 32, G = K * NW * CL;
    const uint32_t SYM = CL > 1 ? MDC_IT_SYM : MD2_IT_SYM;
    const int TSHIFT = CL > 1 ? MDC_IT_TSHIFT : MD2_IT_TSHIFT;
    const uint32_t MAXT = CL > 1 ? (uint32_t)MDC_MAX_TYPES : (uint32_t)MD2_MAX_TYPES;
    const size_t ns = h->desc.size();
    h->md2_desc.assign(ns, Md2Desc());
    h->md2_perm.assign(ns * (size_t)np, MD2_NONE);
    h->md2_eblk.assign(ns * (size_t)G, make_int4(0, 0, 0, 0));
    h->md2_ell_off.clear();
    h->md2_error.clear();
    std::vector<std::vector<uint32_t>> items(ns);
    std::vector<int64_t> nsym(ns, 0), nbond(ns, 0);
    std::vector<std::string> errs(ns);
    auto build_one = [&](size_t s) {
        const StructDesc& d = h->desc[s];
        const int n = d.n_atoms;
        const int32_t* bp = h->bond_ptr.data() + d.bond_ptr_off;
        const uint32_t* bi = h->bond_items.data() + d.bond_item_off;
        const bool a_ok = h->btype_f[d.btype_off].y > 0.f;              // type 0 is an upper bound: block A may be used
        const uint8_t* kindp = h->kind.data() + d.atom_off;
        // ---- orientation: the owner is the end that owns fewer items so far; the other end must have a free inbox slot ----
        struct Own { int32_t other; uint32_t type; int8_t q; };         // q < 0: SYM
        std::vector<std::vector<Own>> rows((size_t)n);
        std::vector<int> own((size_t)n, 0), indeg((size_t)n, 0);
        const int per_cta0 = (n + CL - 1) / CL;                         // atoms per CTA of a cluster (index order)
        for (int a = 0; a < n; a++)
            for (int c = bp[a]; c < bp[a + 1]; c++) {
                const int b = (int)(bi[c] & 0xffffu);
                const uint32_t t = bi[c] >> 16;
                if (b <= a) continue;                                   // every bond once (b == a: a bond to oneself has no force)
                if (t >= MAXT) { errs[s] = "too many distinct bond types in one structure for the MD kernel's item format"; return; }
                nbond[s]++;
                // a frozen particle (the static centre of an Envelope / Damid / Fish restraint, hundreds of bonds in a FISH run)
                // discards its force: the bond lives at the moving end only, without inbox (SYM item); a thread that owned
                // all bonds of such a hub would walk them alone, one dependent load after the other
                const bool fa = kindp[a] == IGMK_KIND_STATIC, fb = kindp[b] == IGMK_KIND_STATIC;
                if (fa || fb) {
                    if (!fa) { rows[a].push_back({b, t, -1}); own[a]++; }
                    if (!fb) { rows[b].push_back({a, t, -1}); own[b]++; }
                    continue;
                }
                const int da = bp[a + 1] - bp[a], db = bp[b + 1] - bp[b];
                int o = (own[a] < own[b] || (own[a] == own[b] && da >= db)) ? a : b, r = o == a ? b : a;
                if (indeg[r] >= MD2_INBOX - 1) std::swap(o, r);
                // k_md2c: a bond between two CTAs of the cluster is evaluated at both ends, which keeps inboxes and mailboxes
                // CTA-local and spares the kernel one of its three cluster barriers per step
                const bool cross = CL > 1 && a / per_cta0 != b / per_cta0;
                if (cross || indeg[r] >= MD2_INBOX - 1) {              // (or both inboxes full): evaluate from both ends
                    rows[a].push_back({b, t, -1}); rows[b].push_back({a, t, -1});
                    own[a]++; own[b]++; nsym[s]++;
                    continue;
                }
                rows[o].push_back({r, t, (int8_t)indeg[r]});
                indeg[r]++; own[o]++;
            }
        // ---- atom order: per CTA by owned items ----
        std::vector<int> order((size_t)n), pos((size_t)n, -1);
        uint16_t* perm = h->md2_perm.data() + s * (size_t)np;
        auto slot_of_group = [&](int g) { const int k = g / NW, r = g % NW; return k * NW + ((k & 1) ? NW - 1 - r : r); };
        const int per_cta = (n + CL - 1) / CL;
        for (int r = 0; r < CL; r++) {
            const int a0 = std::min(n, r * per_cta), a1 = std::min(n, a0 + per_cta), nr = a1 - a0;
            for (int a = a0; a < a1; a++) order[a] = a;
            std::stable_sort(order.begin() + a0, order.begin() + a1, [&](int a, int b) { return own[a] > own[b]; });
            for (int g = 0; g < K * NW; g++) {
                const int gi = slot_of_group(g), k = gi / NW, w = gi % NW;
                for (int l = 0; l < 32; l++) {
                    const int q = g * 32 + l;
                    if (q >= nr) break;
                    const int p = r * npc + k * T + w * 32 + l;
                    perm[p] = (uint16_t)order[a0 + q]; pos[order[a0 + q]] = p;
                }
            }
        }
        Md2Desc& m = h->md2_desc[s];
        m.perm_off = (int32_t)(s * (size_t)np);
        m.ell_desc_off = (int32_t)(s * (size_t)G);
        m.K = K;
        int4* eb = h->md2_eblk.data() + m.ell_desc_off;
        std::vector<uint32_t>& out = items[s];
        for (int gi = 0; gi < G; gi++) {               // blocks in (CTA, slot, warp) order = the order the kernel indexes them
            const int w = gi % NW, p0 = (gi / NW) * T + w * 32;           // first position of the warp-slot ((CTA * K + slot) * T + warp * 32)
            int cols[2] = {0, 0};
            for (int l = 0; l < 32; l++) {
                const int a = perm[p0 + l];
                if (a == MD2_NONE) continue;
                int c[2] = {0, 0};
                for (const Own& it : rows[a]) c[(a_ok && it.type == 0 && it.q >= 0) ? 0 : 1]++;
                cols[0] = std::max(cols[0], c[0]); cols[1] = std::max(cols[1], c[1]);
            }
            int off[2];
            for (int blk = 0; blk < 2; blk++) {
                cols[blk] = (cols[blk] + 3) / 4 * 4;    // the kernel reads 4 columns at a time, unguarded
                off[blk] = (int)out.size();
                out.resize(out.size() + (size_t)cols[blk] * 32, 0u);
                uint32_t* b = out.data() + off[blk];
                // layout [batch of 4 columns][lane][4]: one 16-byte load per lane and batch; padding = an item to oneself
                for (int l = 0; l < 32; l++) {
                    const uint32_t self = ((uint32_t)(p0 + l) << 4) | (blk == 0 ? 15u : SYM);
                    for (int c = 0; c < cols[blk]; c++) b[((c >> 2) * 32 + l) * 4 + (c & 3)] = self;
                }
                for (int l = 0; l < 32; l++) {
                    const int a = perm[p0 + l];
                    if (a == MD2_NONE) continue;
                    int c = 0;
                    for (const Own& it : rows[a]) {
                        if (((a_ok && it.type == 0 && it.q >= 0) ? 0 : 1) != blk) continue;
                        const uint32_t v = ((uint32_t)pos[it.other] << 4) | (it.q >= 0 ? (uint32_t)it.q : SYM) | (it.type << TSHIFT);
                        b[((c >> 2) * 32 + l) * 4 + (c & 3)] = v;
                        c++;
                    }
                }
            }
            eb[gi] = make_int4(off[0], cols[0] / 4, off[1], cols[1] / 4);
        }
    };
    const int nthr = (int)std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), ns));
    std::vector<std::thread> pool;
    for (int t = 0; t < nthr; t++)
        pool.emplace_back([&, t]() { for (size_t s = (size_t)t; s < ns; s += (size_t)nthr) build_one(s); });
    for (auto& th : pool) th.join();
    size_t total = 0;
    h->md2_sym_bonds = 0; h->md2_bonds = 0;
    for (size_t s = 0; s < ns; s++) {
        if (!errs[s].empty() && h->md2_error.empty()) h->md2_error = errs[s];
        h->md2_desc[s].ell_item_off = (int64_t)total; total += items[s].size();
        h->md2_sym_bonds += nsym[s]; h->md2_bonds += nbond[s];
    }
    h->md2_ell_items.resize(total);
    pool.clear();
    for (int t = 0; t < nthr; t++)
        pool.emplace_back([&, t]() {
            for (size_t s = (size_t)t; s < ns; s += (size_t)nthr)
                if (!items[s].empty())
                    memcpy(h->md2_ell_items.data() + h->md2_desc[s].ell_item_off, items[s].data(), items[s].size() * sizeof(uint32_t));
        });
    for (auto& th : pool) th.join();
}

static int upload_dev_rebuild(igmk_handle* h);
static int mdc_resident_clusters(int K, int CL);

static int upload(igmk_handle* h)
{
    if (h->device < 0) return fail("igmk: this handle was created with IGMK_STAGING_ONLY and cannot touch a GPU");
    if (!h->dirty) return 0;
    if (h->dev_built) return upload_dev_rebuild(h);
    h->free_device();
    const size_t na = h->xyzr.size(), ns = h->desc.size();
    CU(dev_alloc_copy(&h->d_desc, h->desc, h->stream));
    CU(dev_alloc_copy(&h->d_xyzr, h->xyzr, h->stream));
    CU(dev_alloc_copy(&h->d_xyzr0, h->xyzr0, h->stream));
    CU(dev_alloc_copy(&h->d_kind, h->kind, h->stream));
    CU(dev_alloc_copy(&h->d_envmask, h->envmask, h->stream));
    CU(dev_alloc_copy(&h->d_bond_ptr, h->bond_ptr, h->stream));
    CU(dev_alloc_copy(&h->d_bond_items, h->bond_items, h->stream));
    CU(dev_alloc_copy(&h->d_btype_f, h->btype_f, h->stream));
    CU(dev_alloc_copy(&h->d_btype_d, h->btype_d, h->stream));
    CU(cudaMalloc((void**)&h->d_nl, std::max<size_t>(na, 1) * h->nl_width * sizeof(uint16_t)));
    CU(cudaMalloc((void**)&h->d_nl_cnt, std::max<size_t>(na, 1) * sizeof(uint16_t)));
    CU(cudaMalloc((void**)&h->d_xd, std::max<size_t>(na, 1) * 3 * sizeof(double)));
    CU(cudaMalloc((void**)&h->d_fd, std::max<size_t>(na, 1) * 3 * sizeof(double)));
    CU(cudaMalloc((void**)&h->d_work, std::max<size_t>(na, 1) * 12 * sizeof(double)));
    CU(cudaMalloc((void**)&h->d_res, std::max<size_t>(ns, 1) * sizeof(igmk_result_t)));
    CU(cudaMalloc((void**)&h->d_flags, std::max<size_t>(ns, 1) * sizeof(int32_t)));
    CU(cudaMemsetAsync(h->d_res, 0, std::max<size_t>(ns, 1) * sizeof(igmk_result_t), h->stream));
    CU(cudaMemsetAsync(h->d_flags, 0, std::max<size_t>(ns, 1) * sizeof(int32_t), h->stream));
    h->md2_K = md2_pick_K(h->max_atoms);
    h->md_big = h->md2_K == 0 || getenv("IGMK_FORCE_BIG") != nullptr;   // the variable routes small batches through md_big.cuh (tests)
    h->mdc_CL = 0;
    {   // more than 4096 atoms: a thread-block cluster per structure (md_cluster.cuh); IGMK_FORCE_CLUSTER=<CTAs> sends smaller
        // structures there too (tests), IGMK_FORCE_BIG keeps the one-CTA global-memory kernel
        int want = 0;
        if (const char* e = getenv("IGMK_FORCE_CLUSTER")) want = atoi(e);
        if (want == 0 && h->md2_K == 0 && !getenv("IGMK_FORCE_BIG") && h->max_atoms <= 65535) {
            // Cheapest of: one CTA per structure with the state in global memory (k_md_big: 29 ns per atom and evaluation,
            // every SM busy once there are 148 structures), or clusters of 16 / 8 / 4 / 2 CTAs run in waves of as many
            // clusters as the device holds at once (the driver's answer: a cluster must sit inside one GPC, so a B200 holds
            // 8 clusters of 16, not 148 / 16 = 9).  Cluster cost per atom and evaluation measured at 29 838 beads
            // (tools/prof_c4.py, round 2): 2.1 ns on 16 CTAs, 3.2 ns on 8; smaller clusters extrapolated with the same slope.
            const long long nsl = (long long)ns;
            double best = (double)((nsl + h->sm_count - 1) / h->sm_count) * 29.0;
            for (int cl = 16; cl >= 2; cl >>= 1) {
                if ((long long)cl * 4 * MD2_T < h->max_atoms) continue;
                int k = 1;
                while (k < 4 && k * MD2_T * cl < h->max_atoms) k *= 2;
                const int resident = mdc_resident_clusters(k, cl);
                if (resident <= 0) continue;
                const double cost = (double)((nsl + resident - 1) / resident) * 2.1 * pow(16.0 / cl, 0.585);
                if (cost < best) { best = cost; want = cl; }
                if (getenv("IGMK_TRACE")) fprintf(stderr, "[igmk trace] %d clusters of %d CTAs (K = %d) fit the device; cost %.1f\n", resident, cl, k, cost);
            }
        }
        if (want == 2 || want == 4 || want == 8 || want == 16) {
            int k = 1;
            while (k < 4 && k * MD2_T * want < h->max_atoms) k *= 2;
            if (k * MD2_T * want >= h->max_atoms && !getenv("IGMK_FORCE_BIG")) { h->mdc_CL = want; h->md2_K = k; h->md_big = false; }
        }
    }
    if (h->md_big) {                                   // > 8192 atoms: state in global memory (md_big.cuh)
        h->md2_K = (h->max_atoms + MD2_T - 1) / MD2_T;
        if ((long long)md_big_smem_bytes(h->md2_K) > (long long)h->smem_optin)
            return fail("igmk: structure too large even for the global-memory MD path (cell grid exceeds shared memory)");
        CU(cudaMalloc((void**)&h->d_big_state, ns * (size_t)h->md2_K * MD2_T * BIG_FLOATS_PER_POS * sizeof(float)));
    }
    if (ctx_is_big<double>(h->max_atoms, false) || ctx_is_big<float>(h->max_atoms, false))
        CU(cudaMalloc((void**)&h->d_big, (na + 32 * ns) * 10 * sizeof(double)));
    if (h->md2_K > 0) {
        if (h->md_big) md2_build(h);
        else {
            md2_build_owner(h, std::max(h->mdc_CL, 1));
            if (!h->md2_error.empty()) return fail("igmk: " + h->md2_error);
        }
        CU(dev_alloc_copy(&h->d_md2_desc, h->md2_desc, h->stream));
        CU(dev_alloc_copy(&h->d_md2_perm, h->md2_perm, h->stream));
        CU(dev_alloc_copy(&h->d_md2_ell_off, h->md2_ell_off, h->stream));
        CU(dev_alloc_copy(&h->d_md2_ell_items, h->md2_ell_items, h->stream));
        CU(cudaMalloc((void**)&h->d_md2_ell_act, std::max<size_t>(h->md2_ell_items.size(), 1) * sizeof(uint32_t)));
        CU(cudaMalloc((void**)&h->d_md2_nl, h->md2_perm.size() * (size_t)h->md2_nlw * sizeof(uint16_t)));
        if (!h->md_big) {
            CU(dev_alloc_copy(&h->d_md2_eblk, h->md2_eblk, h->stream));
            CU(cudaMalloc((void**)&h->d_md2_gf, h->md2_perm.size() * (size_t)MD2_INBOX * sizeof(float4)));
            if (h->mdc_CL > 0) {
                const size_t npos = h->md2_perm.size();
                CU(cudaMalloc((void**)&h->d_mdc_xall, npos * sizeof(float4)));
                CU(cudaMalloc((void**)&h->d_mdc_xsorted, npos * sizeof(float4)));
                CU(cudaMalloc((void**)&h->d_mdc_sorted, npos * sizeof(uint32_t)));
                CU(cudaMalloc((void**)&h->d_mdc_rowlen, npos * sizeof(uint16_t)));
                CU(cudaMalloc((void**)&h->d_mdc_cells, ns * 2 * (size_t)(MDC_MAXC + 1) * sizeof(uint32_t)));
                CU(cudaMalloc((void**)&h->d_mdc_bbox, ns * (size_t)MDC_MAX_CL * 8 * sizeof(float)));
            }
        }
    }
    h->h2d_bytes = ns * sizeof(StructDesc) + 2 * na * sizeof(float4) + na * (sizeof(uint8_t) + sizeof(uint32_t)) +
                   h->bond_ptr.size() * sizeof(int32_t) + h->bond_items.size() * sizeof(uint32_t) +
                   h->btype_f.size() * (sizeof(float2) + sizeof(double2)) +
                   h->md2_desc.size() * sizeof(Md2Desc) + h->md2_perm.size() * sizeof(uint16_t) +
                   h->md2_ell_off.size() * sizeof(int32_t) + h->md2_ell_items.size() * sizeof(uint32_t) +
                   (h->md_big ? 0 : h->md2_eblk.size() * sizeof(int4));
    h->dirty = false;
    return 0;
}

extern "C" int igmk_upload(igmk_handle_t* h)
{
    if (!h) return fail("igmk_upload: NULL handle");
    if (h->desc.empty()) return fail("igmk_upload: nothing submitted");
    if (h->device < 0) return upload(h);
    CU(cudaSetDevice(h->device));
    return upload(h);
}

extern "C" int igmk_layout_stats(igmk_handle_t* h, int32_t* slots_per_thread, int64_t* bond_items, int64_t* ell_items,
                                 double* build_seconds)
{
    if (!h) return fail("igmk_layout_stats: NULL handle");
    if (h->desc.empty()) return fail("igmk_layout_stats: nothing submitted");
    if (h->dev_built) return fail("igmk_layout_stats: not available for a batch assembled on the device");
    int K = md2_pick_K(h->max_atoms);
    if (K == 0) K = (h->max_atoms + MD2_T - 1) / MD2_T;           // global-memory path (md_big.cuh): same layout, more slots
    const int keep_K = h->md2_K;
    h->md2_K = K;
    const auto t0 = std::chrono::steady_clock::now();
    if (md2_pick_K(h->max_atoms) == 0) md2_build(h); else md2_build_owner(h);
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    h->md2_K = keep_K;
    h->dirty = true;                                              // the next upload rebuilds for its own K / path
    if (slots_per_thread) *slots_per_thread = K;
    if (bond_items) *bond_items = (int64_t)h->bond_items.size();
    if (ell_items) *ell_items = (int64_t)h->md2_ell_items.size();
    if (build_seconds) *build_seconds = dt;
    return 0;
}

extern "C" int igmk_slot_sizes(igmk_handle_t* h, int32_t slot, int32_t* n_atoms, int64_t* n_bonds, int64_t* n_env_members)
{
    if (!h) return fail("igmk_slot_sizes: NULL handle");
    if (slot < 0 || slot >= (int)h->desc.size()) return fail("igmk_slot_sizes: bad slot");
    const bool last = slot + 1 >= (int)h->desc.size();
    if (n_atoms) *n_atoms = h->desc[slot].n_atoms;
    if (h->dev_built) {
        if (n_bonds) *n_bonds = h->dev_nb[slot];
        if (n_env_members) *n_env_members = h->dev_env_off.empty() ? 0 : h->dev_env_off.back();
        return 0;
    }
    if (n_bonds) *n_bonds = (last ? (int64_t)h->sub_bonds.size() : h->sub_bond_start[slot + 1]) - h->sub_bond_start[slot];
    if (n_env_members) *n_env_members = (last ? (int64_t)h->sub_env.size() : h->sub_env_start[slot + 1]) - h->sub_env_start[slot];
    return 0;
}

// M10: violation ratios of every submitted restraint at the batch's current coordinates
extern "C" int igmk_violation_ratios(igmk_handle_t* h, int32_t slot, double* bond_ratio, double* env_u)
{
    if (!h) return fail("igmk_violation_ratios: NULL handle");
    if (slot < 0 || slot >= (int)h->desc.size()) return fail("igmk_violation_ratios: bad slot");
    if (h->device < 0) return upload(h);
    CU(cudaSetDevice(h->device));
    if (!h->viol_valid) {
        if (int rc = upload(h)) return rc;
        if (h->dev_built && h->sub_env_start.empty()) {       // envelope members of a device-assembled batch: the template's, per structure
            for (size_t s = 0; s < h->desc.size(); s++) {
                h->sub_env_start.push_back((int64_t)h->sub_env.size());
                for (size_t e = 0; e + 1 < h->dev_env_off.size(); e++)
                    for (int q = h->dev_env_off[e]; q < h->dev_env_off[e + 1]; q++) {
                        EnvMember m;
                        m.atom = h->desc[s].atom_off + h->dev_env_mem[q]; m.pad = 0;
                        m.a = h->dev_env_abck[4 * e]; m.b = h->dev_env_abck[4 * e + 1]; m.c = h->dev_env_abck[4 * e + 2]; m.k = h->dev_env_abck[4 * e + 3];
                        h->sub_env.push_back(m);
                    }
            }
        }
        const int64_t nb = h->dev_built ? h->dev_bonds : (int64_t)h->sub_bonds.size(), ne = (int64_t)h->sub_env.size();
        h->viol_bond.assign((size_t)nb, 0.0); h->viol_env.assign((size_t)ne, 0.0);
        SubBond* d_b = nullptr; int32_t* d_o = nullptr; EnvMember* d_e = nullptr; double* d_r = nullptr;
        cudaError_t e = cudaSuccess;
        if (h->dev_built) {                                   // the bond list never left the device
            int64_t* d_start = nullptr;
            e = cudaMalloc((void**)&d_b, (size_t)std::max<int64_t>(nb, 1) * sizeof(SubBond));
            if (e == cudaSuccess) e = cudaMalloc((void**)&d_o, (size_t)std::max<int64_t>(nb, 1) * sizeof(int32_t));
            if (e == cudaSuccess) e = dev_alloc_copy(&d_start, h->sub_bond_start, h->stream);
            if (e == cudaSuccess && nb > 0) {
                k_asm_sub_bonds<<<(int)std::min<int64_t>((nb + 255) / 256, 148 * 16), 256, 0, h->stream>>>(
                    nb, h->d_asm_bl, h->d_asm_bl + nb, h->d_asm_bl + 2 * nb, h->d_asm_key_r0, h->d_asm_key_k, h->d_asm_key_style,
                    d_start, (int32_t)h->desc.size(), h->d_desc, d_b, d_o);
                e = cudaGetLastError();
                if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
            }
            cudaFree(d_start);
        } else {
            e = dev_alloc_copy(&d_b, h->sub_bonds, h->stream);
            if (e == cudaSuccess) e = dev_alloc_copy(&d_o, h->sub_bond_atom_off, h->stream);
        }
        if (e == cudaSuccess) e = dev_alloc_copy(&d_e, h->sub_env, h->stream);
        if (e == cudaSuccess) e = cudaMalloc((void**)&d_r, (size_t)std::max<int64_t>(std::max(nb, ne), 1) * sizeof(double));
        if (e == cudaSuccess && nb > 0) {
            k_bond_violations<<<(int)std::min<int64_t>((nb + 255) / 256, 148 * 16), 256, 0, h->stream>>>(h->d_xyzr, d_b, d_o, nb, d_r);
            e = cudaGetLastError();
            if (e == cudaSuccess) e = copy_on(h->stream, h->viol_bond.data(), d_r, (size_t)nb * sizeof(double), cudaMemcpyDeviceToHost);
        }
        if (e == cudaSuccess && ne > 0) {
            k_env_violations<<<(int)std::min<int64_t>((ne + 255) / 256, 148 * 16), 256, 0, h->stream>>>(h->d_xyzr, d_e, ne, d_r);
            e = cudaGetLastError();
            if (e == cudaSuccess) e = copy_on(h->stream, h->viol_env.data(), d_r, (size_t)ne * sizeof(double), cudaMemcpyDeviceToHost);
        }
        cudaFree(d_b); cudaFree(d_o); cudaFree(d_e); cudaFree(d_r);
        CU(e);
        h->viol_valid = true;
    }
    const int64_t b0 = h->sub_bond_start[slot], b1 = slot + 1 < (int)h->desc.size() ? h->sub_bond_start[slot + 1]
                                                   : (h->dev_built ? h->dev_bonds : (int64_t)h->sub_bonds.size());
    const int64_t e0 = h->sub_env_start[slot], e1 = slot + 1 < (int)h->desc.size() ? h->sub_env_start[slot + 1] : (int64_t)h->sub_env.size();
    if (bond_ratio && b1 > b0) memcpy(bond_ratio, h->viol_bond.data() + b0, (size_t)(b1 - b0) * sizeof(double));
    if (env_u && e1 > e0) memcpy(env_u, h->viol_env.data() + e0, (size_t)(e1 - e0) * sizeof(double));
    return 0;
}

extern "C" int igmk_last_timings(igmk_handle_t* h, double* md_ms, double* cg_ms, int64_t* h2d_bytes)
{
    if (!h) return fail("igmk_last_timings: NULL handle");
    if (md_ms) *md_ms = h->md_ms;
    if (cg_ms) *cg_ms = h->cg_ms;
    if (h2d_bytes) *h2d_bytes = (int64_t)h->h2d_bytes;
    return 0;
}

extern "C" int igmk_reset_coordinates(igmk_handle_t* h)
{
    if (!h) return fail("igmk_reset_coordinates: NULL handle");
    h->viol_valid = false;
    if (!h->dev_built) h->xyzr = h->xyzr0;       // (a device-assembled batch keeps its submitted coordinates on the device only)
    if (!h->dirty && h->d_xyzr) {
        CU(cudaSetDevice(h->device));
        CU(copy_on(h->stream, h->d_xyzr, h->d_xyzr0, h->xyzr.size() * sizeof(float4), cudaMemcpyDeviceToDevice));
    }
    return 0;
}

static Batch make_batch(igmk_handle* h)
{
    Batch B;
    B.n_struct = (int32_t)h->desc.size(); B.max_atoms = h->max_atoms;
    B.desc = h->d_desc; B.xyzr = h->d_xyzr; B.kind = h->d_kind; B.envmask = h->d_envmask;
    B.bond_ptr = h->d_bond_ptr; B.bond_items = h->d_bond_items; B.btype_f = h->d_btype_f; B.btype_d = h->d_btype_d;
    B.nl = h->d_nl; B.nl_width = h->nl_width; B.nl_cnt = h->d_nl_cnt; B.xd = h->d_xd; B.fd = h->d_fd; B.work = h->d_work; B.big = h->d_big;
    B.res = h->d_res; B.flags = h->d_flags;
    return B;
}

static int collect(igmk_handle* h, const char* what)
{
    const size_t ns = h->desc.size();
    h->res.resize(ns); h->flags.resize(ns);
    CU(copy_on(h->stream, h->flags.data(), h->d_flags, ns * sizeof(int32_t), cudaMemcpyDeviceToHost));
    for (size_t s = 0; s < ns; s++) {
        if (h->flags[s] & 1) {
            char msg[256];
            snprintf(msg, sizeof msg, "%s: neighbour list overflow (> %d neighbours per bead) in structure %lld",
                     what, h->nl_width, (long long)h->struct_id[s]);
            fail(msg);
            return IGMK_NL_OVERFLOW;      // host staging still holds the submitted coordinates
        }
    }
    CU(copy_on(h->stream, h->res.data(), h->d_res, ns * sizeof(igmk_result_t), cudaMemcpyDeviceToHost));
    CU(copy_on(h->stream, h->xyzr.data(), h->d_xyzr, h->xyzr.size() * sizeof(float4), cudaMemcpyDeviceToHost));
    for (size_t s = 0; s < ns; s++) h->res[s].status = (h->flags[s] & 2) ? 1 : 0;
    return 0;
}

// widen the neighbour table after an overflow; false when it already holds every atom
static bool widen_neighbour_table(igmk_handle* h)
{
    if (h->nl_width >= h->max_atoms) return false;
    h->nl_width = std::min(h->nl_width * 2, (h->max_atoms + 7) / 8 * 8);
    h->dirty = true;
    return true;
}

template <int K>
static int launch_md2_k(igmk_handle* h, const Batch& B, const MdParams& P, int ns)
{
    const size_t sm = md2_smem_bytes(K);
    CU(cudaFuncSetAttribute(k_md2<K, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    CU(cudaFuncSetAttribute(k_md2<K, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    Md2View V;
    V.perm = h->d_md2_perm; V.ell_off = h->d_md2_ell_off; V.ell_items = h->d_md2_ell_items;
    V.nl = h->d_md2_nl; V.nl_width = h->md2_nlw; V.ell_act = h->d_md2_ell_act;
    V.eblk = h->d_md2_eblk; V.gf = h->d_md2_gf; V.dbg_forces = h->d_md2_dbg;
    V.prof = nullptr;
    const bool prof = getenv("IGMK_PROF") != nullptr && P.debug_stage < 0;
    if (prof) { CU(cudaMalloc((void**)&V.prof, (size_t)ns * 16 * sizeof(long long))); CU(cudaMemsetAsync(V.prof, 0, (size_t)ns * 16 * sizeof(long long), h->stream)); }
    if (prof) k_md2<K, true><<<ns, MD2S_T, sm, h->stream>>>(B, V, h->d_md2_desc, P);
    else k_md2<K, false><<<ns, MD2S_T, sm, h->stream>>>(B, V, h->d_md2_desc, P);
    CU(cudaGetLastError());
    if (prof) {                  // diagnostics only: per-phase SM cycles of thread 0, averaged over the structures
        std::vector<long long> t((size_t)ns * 16);
        CU(copy_on(h->stream, t.data(), V.prof, t.size() * sizeof(long long), cudaMemcpyDeviceToHost));
        cudaFree(V.prof);
        double a[16] = {0};
        for (int s = 0; s < ns; s++) for (int k = 0; k < 16; k++) a[k] += (double)t[16 * (size_t)s + k] / ns;
        fprintf(stderr, "[igmk prof] k_md2 mean cycles per structure (thread 0): force evaluations %.3g (pair loops %.3g, bond block A %.3g, "
                "barrier + receive %.3g; loops before the barrier: slowest warp %.3g, mean warp %.3g); rebuilds %.3g (of which search %.3g: slowest warp %.3g, "
                "mean warp %.3g), %.1f rebuilds; bond batches walked by thread 0: %.0f candidate, %.0f full; neighbour batches %.0f\n",
                a[0], a[5], a[6], a[4], a[9], a[10], a[2], a[1], a[7], a[8], a[3], a[11], a[12], a[13]);
    }
    return 0;
}

template <int K, int CL>
static int launch_mdc_k(igmk_handle* h, const Batch& B, const MdParams& P, int ns)
{
    const size_t sm = mdc_smem_bytes(K);
    auto kern = k_md2c<K, CL, false>;
    CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    if (CL > 8) CU(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    MdcView V;
    memset(&V, 0, sizeof V);
    V.perm = h->d_md2_perm; V.ell_items = h->d_md2_ell_items; V.ell_act = h->d_md2_ell_act; V.eblk = h->d_md2_eblk;
    V.nl = h->d_md2_nl; V.nl_width = h->md2_nlw; V.gf = h->d_md2_gf;
    V.xall = h->d_mdc_xall; V.xsorted = h->d_mdc_xsorted; V.sorted = h->d_mdc_sorted; V.cells = h->d_mdc_cells;
    V.rowlen = h->d_mdc_rowlen; V.bbox = h->d_mdc_bbox; V.dbg_forces = h->d_md2_dbg;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3((unsigned)(ns * CL)); cfg.blockDim = dim3(MD2_T); cfg.dynamicSmemBytes = sm; cfg.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    const Md2Desc* md = h->d_md2_desc;
    CU(cudaLaunchKernelEx(&cfg, kern, B, V, md, P));
    return 0;
}

template <int K, int CL>
static int mdc_resident_clusters_k()
{
    const size_t sm = mdc_smem_bytes(K);
    auto kern = k_md2c<K, CL, false>;
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm) != cudaSuccess) { cudaGetLastError(); return 0; }
    if (CL > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) { cudaGetLastError(); return 0; }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3((unsigned)CL); cfg.blockDim = dim3(MD2_T); cfg.dynamicSmemBytes = sm;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// clusters of CL CTAs of k_md2c<K, CL> the device holds at once (0: no such kernel)
static int mdc_resident_clusters(int K, int CL)
{
#define MDC_CASE(k, c) if (K == k && CL == c) return mdc_resident_clusters_k<k, c>()
    MDC_CASE(1, 2); MDC_CASE(2, 2); MDC_CASE(4, 2); MDC_CASE(1, 4); MDC_CASE(2, 4); MDC_CASE(4, 4);
    MDC_CASE(1, 8); MDC_CASE(2, 8); MDC_CASE(4, 8); MDC_CASE(1, 16); MDC_CASE(2, 16); MDC_CASE(4, 16);
#undef MDC_CASE
    return 0;
}

static int launch_mdc(igmk_handle* h, const Batch& B, const MdParams& P, int ns)
{
    const int K = h->md2_K, CL = h->mdc_CL;
#define MDC_CASE(k, c) if (K == k && CL == c) return launch_mdc_k<k, c>(h, B, P, ns)
    MDC_CASE(1, 2); MDC_CASE(2, 2); MDC_CASE(4, 2); MDC_CASE(1, 4); MDC_CASE(2, 4); MDC_CASE(4, 4);
    MDC_CASE(1, 8); MDC_CASE(2, 8); MDC_CASE(4, 8); MDC_CASE(1, 16); MDC_CASE(2, 16); MDC_CASE(4, 16);
#undef MDC_CASE
    return fail("igmk_run: no cluster MD kernel for this structure size / cluster size");
}

static int launch_md2(igmk_handle* h, const Batch& B, const MdParams& P, int ns)
{
    if (h->mdc_CL > 0) return launch_mdc(h, B, P, ns);
    if (h->md_big) {
        const size_t sm = md_big_smem_bytes(h->md2_K);
        CU(cudaFuncSetAttribute(k_md_big, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        Md2View V;
        memset(&V, 0, sizeof V);
        V.perm = h->d_md2_perm; V.ell_off = h->d_md2_ell_off; V.ell_items = h->d_md2_ell_items;
        V.nl = h->d_md2_nl; V.nl_width = h->md2_nlw; V.prof = nullptr; V.ell_act = h->d_md2_ell_act;
        V.dbg_forces = h->d_md2_dbg;
        k_md_big<<<ns, MD2_T, sm, h->stream>>>(B, V, h->d_md2_desc, P, h->d_big_state);
        CU(cudaGetLastError());
        return 0;
    }
    switch (h->md2_K) {
        case 1: return launch_md2_k<1>(h, B, P, ns);
        case 2: return launch_md2_k<2>(h, B, P, ns);
        case 3: return launch_md2_k<3>(h, B, P, ns);
        case 4: return launch_md2_k<4>(h, B, P, ns);
        case 5: return launch_md2_k<5>(h, B, P, ns);
    }
    return fail("igmk_run: no MD kernel for this structure size");
}

extern "C" int igmk_run(igmk_handle_t* h, const igmk_protocol_t* prot)
{
    if (!h || !prot) return fail("igmk_run: NULL argument");
    if (h->desc.empty()) return fail("igmk_run: nothing submitted");
    if (prot->n_stage < 0 || prot->n_stage > IGMK_MAX_STAGES) return fail("igmk_run: bad n_stage");
    if (!(prot->timestep > 0.0)) return fail("igmk_run: timestep must be > 0");
    if (h->device < 0) return upload(h);
    CU(cudaSetDevice(h->device));
    const int ns = (int)h->desc.size();
    const size_t sm_cg = ctx_smem_bytes<double>(h->max_atoms, false);
    CU(cudaFuncSetAttribute(k_cg, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_cg));
    float ms = 0.f;
    const bool trace = getenv("IGMK_TRACE") != nullptr;     // wall-clock marks of this call on stderr (pipeline diagnostics)
    auto now_ms = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_enter = now_ms();
    for (;;) {
        if (int rc = upload(h)) return rc;
        const double t_up = now_ms();
        Batch B = make_batch(h);
        CU(cudaMemsetAsync(h->d_flags, 0, ns * sizeof(int32_t), h->stream));
        cudaEvent_t e0, e1, e2;
        CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1)); CU(cudaEventCreate(&e2));
        CU(cudaEventRecord(e0, h->stream));
        MdParams P; P.prot = *prot; P.skin_factor = h->md_big ? h->skin_factor : (h->mdc_CL > 0 ? h->skin_cluster : h->skin_inner);
        P.debug_stage = -1; P.general_pairs = getenv("IGMK_GENERAL_PAIRS") != nullptr;
        if (int rc = launch_md2(h, B, P, ns)) return rc;
        CU(cudaEventRecord(e1, h->stream));
        CU(cudaStreamWaitEvent(h->stream_hi, e1, 0));
        k_cg<<<ns, MD_THREADS, sm_cg, h->stream_hi>>>(B, *prot);
        CU(cudaGetLastError());
        CU(cudaEventRecord(e2, h->stream_hi));
        CU(cudaEventSynchronize(e2));
        float ms_md = 0.f, ms_cg = 0.f;
        CU(cudaEventElapsedTime(&ms_md, e0, e1));
        CU(cudaEventElapsedTime(&ms_cg, e1, e2));
        ms = ms_md + ms_cg; h->md_ms = ms_md; h->cg_ms = ms_cg;
        if (trace)
            fprintf(stderr, "[igmk trace] run of %d structures (first id %lld): enter %.1f, uploaded +%.1f, device done +%.1f (md %.1f, cg %.1f)\n",
                    ns, (long long)h->struct_id[0], t_enter, t_up - t_enter, now_ms() - t_enter, ms_md, ms_cg);
        cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
        const int rc = collect(h, "igmk_run");
        if (rc == IGMK_NL_OVERFLOW) {                                  // either table may have overflowed
            const int cap = (h->max_atoms + 7) / 8 * 8;
            const bool w1 = widen_neighbour_table(h);
            const bool w2 = h->md2_nlw < cap;
            if (w2) { h->md2_nlw = std::min(h->md2_nlw * 2, cap); h->dirty = true; }
            if (w1 || w2) continue;
        }
        if (rc) return 1;
        break;
    }
    h->viol_valid = false;
    for (auto& r : h->res) r.md_seconds = ms * 1e-3;
    h->forces.clear();                 // copied on demand by igmk_fetch_forces
    h->have_forces = true;
    return 0;
}

// Forces as the MD kernel itself computes them (float32, its own lists and approximations), for the parity test of
// that routine: run setup of annealing stage `stage` (its pair and envelope scales), no time step taken.
extern "C" int igmk_md_forces(igmk_handle_t* h, const igmk_protocol_t* prot, int32_t stage)
{
    if (!h || !prot) return fail("igmk_md_forces: NULL argument");
    if (h->desc.empty()) return fail("igmk_md_forces: nothing submitted");
    if (stage < 0 || stage >= prot->n_stage) return fail("igmk_md_forces: no such stage");
    if (h->device < 0) return upload(h);
    CU(cudaSetDevice(h->device));
    const int ns = (int)h->desc.size();
    for (;;) {
        if (int rc = upload(h)) return rc;
        if (!h->d_md2_dbg) CU(cudaMalloc((void**)&h->d_md2_dbg, h->xyzr.size() * 3 * sizeof(double)));
        CU(cudaMemsetAsync(h->d_md2_dbg, 0, h->xyzr.size() * 3 * sizeof(double), h->stream));
        Batch B = make_batch(h);
        CU(cudaMemsetAsync(h->d_flags, 0, ns * sizeof(int32_t), h->stream));
        MdParams P; P.prot = *prot; P.skin_factor = h->mdc_CL > 0 ? h->skin_cluster : h->skin_inner; P.debug_stage = stage; P.general_pairs = getenv("IGMK_GENERAL_PAIRS") != nullptr;
        if (int rc = launch_md2(h, B, P, ns)) return rc;
        CU(cudaStreamSynchronize(h->stream));
        const int rc = collect(h, "igmk_md_forces");
        if (rc == IGMK_NL_OVERFLOW) {
            const int cap = (h->max_atoms + 7) / 8 * 8;
            bool w = false;
            if (h->md2_nlw < cap) { h->md2_nlw = std::min(h->md2_nlw * 2, cap); w = true; }
            if (w) { h->dirty = true; continue; }
        }
        if (rc) return 1;
        break;
    }
    h->forces.resize(h->xyzr.size() * 3);
    CU(copy_on(h->stream, h->forces.data(), h->d_md2_dbg, h->forces.size() * sizeof(double), cudaMemcpyDeviceToHost));
    h->have_forces = true;
    return 0;
}

template <typename T>
static int single_point_impl(igmk_handle* h, double ev_scale, double env_scale, double bond_prefactor)
{
    if (!h) return fail("igmk_single_point: NULL handle");
    if (h->desc.empty()) return fail("igmk_single_point: nothing submitted");
    if (h->device < 0) return upload(h);
    CU(cudaSetDevice(h->device));
    const size_t sm = ctx_smem_bytes<T>(h->max_atoms, false);
    CU(cudaFuncSetAttribute(k_single_point<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
    for (;;) {
        if (int rc = upload(h)) return rc;
        Batch B = make_batch(h);
        CU(cudaMemsetAsync(h->d_flags, 0, h->desc.size() * sizeof(int32_t), h->stream));
        k_single_point<T><<<(int)h->desc.size(), MD_THREADS, sm, h->stream>>>(B, ev_scale, env_scale, bond_prefactor);
        CU(cudaGetLastError());
        CU(cudaStreamSynchronize(h->stream));
        const int rc = collect(h, "igmk_single_point");
        if (rc == IGMK_NL_OVERFLOW && widen_neighbour_table(h)) continue;
        if (rc) return 1;
        break;
    }
    h->forces.resize(h->xyzr.size() * 3);
    CU(copy_on(h->stream, h->forces.data(), h->d_fd, h->forces.size() * sizeof(double), cudaMemcpyDeviceToHost));
    h->have_forces = true;
    return 0;
}

extern "C" int igmk_single_point(igmk_handle_t* h, double ev_scale, double env_scale, double bond_prefactor)
{
    return single_point_impl<double>(h, ev_scale, env_scale, bond_prefactor);
}
extern "C" int igmk_single_point_f32(igmk_handle_t* h, double ev_scale, double env_scale, double bond_prefactor)
{
    return single_point_impl<float>(h, ev_scale, env_scale, bond_prefactor);
}

extern "C" int igmk_fetch(igmk_handle_t* h, int32_t slot, float* xyz_out, igmk_result_t* res)
{
    if (!h) return fail("igmk_fetch: NULL handle");
    if (slot < 0 || slot >= (int)h->desc.size()) return fail("igmk_fetch: bad slot");
    if (h->res.size() != h->desc.size()) return fail("igmk_fetch: nothing has been run");
    const StructDesc& d = h->desc[slot];
    if (xyz_out)
        for (int i = 0; i < d.n_atoms; i++) {
            const float4 p = h->xyzr[d.atom_off + i];
            xyz_out[3 * i] = p.x; xyz_out[3 * i + 1] = p.y; xyz_out[3 * i + 2] = p.z;
        }
    if (res) *res = h->res[slot];
    return 0;
}

extern "C" int igmk_fetch_population(igmk_handle_t* h, int32_t n_bead, int32_t n_struct_total, int32_t s_begin, float* crd,
                                     igmk_result_t* res)
{
    if (!h || !crd) return fail("igmk_fetch_population: NULL argument");
    const int ns = (int)h->desc.size();
    if (h->res.size() != h->desc.size()) return fail("igmk_fetch_population: nothing has been run");
    if (n_bead < 1 || s_begin < 0 || s_begin + ns > n_struct_total) return fail("igmk_fetch_population: bad structure range");
    for (int s = 0; s < ns; s++)
        if (h->desc[s].n_atoms < n_bead) return fail("igmk_fetch_population: a structure has fewer than n_bead atoms");
    for (int a = 0; a < n_bead; a++) {                      // bead-major like the .hss dataset: contiguous runs of ns structures
        float* dst = crd + ((size_t)a * n_struct_total + s_begin) * 3;
        for (int s = 0; s < ns; s++) {
            const float4 p = h->xyzr[h->desc[s].atom_off + a];
            dst[3 * s] = p.x; dst[3 * s + 1] = p.y; dst[3 * s + 2] = p.z;
        }
    }
    if (res) memcpy(res, h->res.data(), (size_t)ns * sizeof(igmk_result_t));
    return 0;
}

extern "C" int igmk_fetch_forces(igmk_handle_t* h, int32_t slot, double* f_out)
{
    if (!h || !f_out) return fail("igmk_fetch_forces: NULL argument");
    if (slot < 0 || slot >= (int)h->desc.size()) return fail("igmk_fetch_forces: bad slot");
    if (!h->have_forces) return fail("igmk_fetch_forces: no forces available");
    if (h->forces.size() != h->xyzr.size() * 3) {
        CU(cudaSetDevice(h->device));
        h->forces.resize(h->xyzr.size() * 3);
        CU(copy_on(h->stream, h->forces.data(), h->d_fd, h->forces.size() * sizeof(double), cudaMemcpyDeviceToHost));
    }
    const StructDesc& d = h->desc[slot];
    memcpy(f_out, h->forces.data() + (size_t)d.atom_off * 3, (size_t)d.n_atoms * 3 * sizeof(double));
    return 0;
}

// =========================================================================================
// A-step
template <typename T>
struct DevBuf {
    T* p = nullptr; size_t cap = 0;
    cudaError_t reserve(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc((void**)&p, n * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct igmk_pop {
    int device = 0;
    int smem_optin = 0;
    int sm_count = 0;
    int32_t n_bead = 0, n_struct = 0, n_hap = 0;
    int max_copies = 1;
    float* d_crd = nullptr; bool own_crd = false;
    float* d_radii = nullptr; int32_t* d_chrom = nullptr;
    int32_t* d_copy_off = nullptr; int32_t* d_copy_beads = nullptr;
    std::vector<int32_t> copy_off, copy_beads, chrom;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float last_ms = 0.f;
    // staged pair batch (activation distances)
    int64_t n_pairs = 0;
    std::vector<int32_t> h_pi, h_pj;
    DevBuf<int32_t> d_pi, d_pj; DevBuf<double> d_pw, d_pl;
    DevBuf<float> d_dist, d_prob; DevBuf<uint8_t> d_valid; DevBuf<double> d_den; DevBuf<PairMeta> d_meta;
    std::vector<float> radii;
    bool actdist_done = false;
    // staged actdist rows (assignment)
    int64_t n_rows = 0;
    DevBuf<int32_t> d_row, d_col; DevBuf<float> d_rdist; DevBuf<uint32_t> d_bits;
    bool assign_done = false;
    // bound distance of every staged row as an index into a table of distinct values (igmk_pop_set_row_bounds), for the
    // device-side hand-off to the M-step
    DevBuf<int32_t> d_rkey; std::vector<double> key_r0; double row_k = 0.0; int64_t bounds_rows = -1;

    PopView view() const
    {
        PopView v;
        v.n_bead = n_bead; v.n_struct = n_struct; v.n_hap = n_hap;
        v.crd = d_crd; v.radii = d_radii; v.chrom = d_chrom; v.copy_off = d_copy_off; v.copy_beads = d_copy_beads;
        return v;
    }
    void free_all()
    {
        if (own_crd) cudaFree(d_crd);
        cudaFree(d_radii); cudaFree(d_chrom); cudaFree(d_copy_off); cudaFree(d_copy_beads);
        d_pi.release(); d_pj.release(); d_pw.release(); d_pl.release(); d_dist.release(); d_prob.release();
        d_valid.release(); d_den.release(); d_meta.release(); d_row.release(); d_col.release(); d_rdist.release(); d_bits.release();
        d_rkey.release();
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (stream) cudaStreamDestroy(stream);
    }
};

extern "C" int igmk_pop_create(int device, int32_t n_bead, int32_t n_struct, int32_t n_hap,
                               const float* radii, const int32_t* chrom_bead,
                               const int32_t* copy_off, const int32_t* copy_beads, igmk_pop_t** out)
{
    if (!out || !radii || !chrom_bead || !copy_off || !copy_beads) return fail("igmk_pop_create: NULL argument");
    if (n_bead <= 0 || n_struct <= 0 || n_hap <= 0 || n_hap > n_bead) return fail("igmk_pop_create: bad sizes");
    int n = 0;
    CU(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return fail("igmk_pop_create: no such CUDA device");
    CU(cudaSetDevice(device));
    if (copy_off[0] != 0) return fail("igmk_pop_create: copy_off[0] must be 0");
    int max_copies = 1;
    for (int i = 0; i < n_hap; i++) {
        int c = copy_off[i + 1] - copy_off[i];
        if (c < 1 || c > AD_MAX_COPIES) return fail("igmk_pop_create: every haploid bin needs 1..4 copies");
        max_copies = std::max(max_copies, c);
    }
    for (int q = 0; q < copy_off[n_hap]; q++)
        if (copy_beads[q] < 0 || copy_beads[q] >= n_bead) return fail("igmk_pop_create: copy bead out of range");
    igmk_pop* p = new igmk_pop();
    p->device = device; p->n_bead = n_bead; p->n_struct = n_struct; p->n_hap = n_hap; p->max_copies = max_copies;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    p->smem_optin = (int)prop.sharedMemPerBlockOptin;
    p->sm_count = prop.multiProcessorCount;
    p->copy_off.assign(copy_off, copy_off + n_hap + 1);
    p->copy_beads.assign(copy_beads, copy_beads + copy_off[n_hap]);
    p->chrom.assign(chrom_bead, chrom_bead + n_bead);
    p->radii.assign(radii, radii + n_bead);
    CU(cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&p->ev0)); CU(cudaEventCreate(&p->ev1));
    p->d_crd = nullptr; p->own_crd = false;      // the slab is allocated by the first igmk_pop_upload, or attached by the caller
    CU(cudaMalloc((void**)&p->d_radii, (size_t)n_bead * sizeof(float)));
    CU(cudaMalloc((void**)&p->d_chrom, (size_t)n_bead * sizeof(int32_t)));
    CU(cudaMalloc((void**)&p->d_copy_off, (size_t)(n_hap + 1) * sizeof(int32_t)));
    CU(cudaMalloc((void**)&p->d_copy_beads, p->copy_beads.size() * sizeof(int32_t)));
    CU(cudaMemcpy(p->d_radii, radii, (size_t)n_bead * sizeof(float), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p->d_chrom, chrom_bead, (size_t)n_bead * sizeof(int32_t), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p->d_copy_off, copy_off, (size_t)(n_hap + 1) * sizeof(int32_t), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(p->d_copy_beads, p->copy_beads.data(), p->copy_beads.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    *out = p;
    return 0;
}

extern "C" int igmk_pop_destroy(igmk_pop_t* p)
{
    if (!p) return 0;
    cudaSetDevice(p->device);
    p->free_all();
    delete p;
    return 0;
}

extern "C" int igmk_pop_upload(igmk_pop_t* p, const float* crd)
{
    if (!p || !crd) return fail("igmk_pop_upload: NULL argument");
    CU(cudaSetDevice(p->device));
    if (!p->d_crd || !p->own_crd) {
        p->d_crd = nullptr; p->own_crd = false;
        CU(cudaMalloc((void**)&p->d_crd, (size_t)p->n_bead * p->n_struct * 3 * sizeof(float)));
        p->own_crd = true;
    }
    CU(cudaMemcpyAsync(p->d_crd, crd, (size_t)p->n_bead * p->n_struct * 3 * sizeof(float), cudaMemcpyHostToDevice, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    return 0;
}

extern "C" int igmk_pop_download(igmk_pop_t* p, float* crd)
{
    if (!p || !crd) return fail("igmk_pop_download: NULL argument");
    if (!p->d_crd) return fail("igmk_pop_download: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    CU(cudaMemcpyAsync(crd, p->d_crd, (size_t)p->n_bead * p->n_struct * 3 * sizeof(float), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    return 0;
}

extern "C" int igmk_pop_coords_devptr(igmk_pop_t* p, void** dptr)
{
    if (!p || !dptr) return fail("igmk_pop_coords_devptr: NULL argument");
    *dptr = p->d_crd;
    return 0;
}

extern "C" int igmk_pop_attach_coords(igmk_pop_t* p, void* dptr)
{
    if (!p || !dptr) return fail("igmk_pop_attach_coords: NULL argument");
    CU(cudaSetDevice(p->device));
    cudaPointerAttributes a;
    CU(cudaPointerGetAttributes(&a, dptr));
    if (a.type != cudaMemoryTypeDevice || a.device != p->device)
        return fail("igmk_pop_attach_coords: not a device pointer of this population's GPU");
    // the slab was produced on a stream of the caller (a torch tensor after an NCCL all-gather, a copy): the kernels of
    // this population run on its own non-blocking stream, which is not ordered after that work
    CU(cudaDeviceSynchronize());
    if (p->own_crd) cudaFree(p->d_crd);
    p->d_crd = (float*)dptr; p->own_crd = false;
    return 0;
}

extern "C" double igmk_pop_last_kernel_ms(igmk_pop_t* p) { return p ? (double)p->last_ms : 0.0; }

extern "C" int igmk_pop_set_pairs(igmk_pop_t* p, int64_t n_pairs, const int32_t* pair_i, const int32_t* pair_j,
                                  const double* pwish, const double* plast)
{
    if (!p) return fail("igmk_pop_set_pairs: NULL handle");
    if (n_pairs < 0) return fail("igmk_pop_set_pairs: negative pair count");
    p->n_pairs = 0; p->actdist_done = false;
    if (n_pairs == 0) return 0;
    if (!pair_i || !pair_j || !pwish || !plast) return fail("igmk_pop_set_pairs: NULL argument");
    for (int64_t k = 0; k < n_pairs; k++)
        if (pair_i[k] < 0 || pair_i[k] >= p->n_hap || pair_j[k] < 0 || pair_j[k] >= p->n_hap)
            return fail("igmk_pop_set_pairs: haploid index out of range");
    CU(cudaSetDevice(p->device));
    CU(p->d_pi.reserve(n_pairs)); CU(p->d_pj.reserve(n_pairs)); CU(p->d_pw.reserve(n_pairs)); CU(p->d_pl.reserve(n_pairs));
    CU(p->d_dist.reserve(n_pairs)); CU(p->d_prob.reserve(n_pairs)); CU(p->d_valid.reserve(n_pairs));
    CU(cudaMemcpyAsync(p->d_pi.p, pair_i, n_pairs * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_pj.p, pair_j, n_pairs * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_pw.p, pwish, n_pairs * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_pl.p, plast, n_pairs * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    p->h_pi.assign(pair_i, pair_i + n_pairs); p->h_pj.assign(pair_j, pair_j + n_pairs);
    CU(cudaStreamSynchronize(p->stream));
    p->n_pairs = n_pairs;
    return 0;
}

extern "C" int igmk_pop_actdist_launch(igmk_pop_t* p, double contact_range)
{
    if (!p) return fail("igmk_pop_actdist_launch: NULL handle");
    p->last_ms = 0.f;
    if (p->n_pairs == 0) { p->actdist_done = true; return 0; }
    if (!p->d_crd) return fail("igmk_pop_actdist_launch: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const int spt = (p->n_struct + AD_THREADS - 1) / AD_THREADS;       // structures per thread
    if (spt <= 8 && p->max_copies * p->max_copies <= 4 && !getenv("IGMK_ACTDIST_SMEM")) {   // the kernel holds 4 combinations
        // register-resident selection (up to 2048 structures)
        void (*kern)(PopView, int64_t, const PairMeta*, float*, float*, uint8_t*) =
            spt <= 1 ? k_actdist_reg<1, 4> : spt <= 2 ? k_actdist_reg<2, 4> : spt <= 4 ? k_actdist_reg<4, 4> : k_actdist_reg<8, 3>;   // measured: 16 keys/thread want 4 CTAs per SM, 32 keys 3
        if (const char* e = getenv("IGMK_ACTDIST_MINB")) {       // resident CTAs per SM the kernel is compiled for (tuning aid)
            const int m = atoi(e);
            if (spt > 2 && spt <= 4) kern = m == 3 ? k_actdist_reg<4, 3> : m == 2 ? k_actdist_reg<4, 2> : kern;
            if (spt > 4) kern = m == 4 ? k_actdist_reg<8, 4> : m == 2 ? k_actdist_reg<8, 2> : kern;
        }
        int per_sm = 1;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, AD_THREADS, 0));
        if (per_sm < 1) per_sm = 1;
        const int64_t grid = std::min<int64_t>(p->n_pairs, (int64_t)p->sm_count * per_sm);
        CU(p->d_meta.reserve((size_t)p->n_pairs));
        CU(cudaEventRecord(p->ev0, p->stream));
        k_actdist_meta<<<(int)std::min<int64_t>((p->n_pairs + 255) / 256, (int64_t)p->sm_count * 8), 256, 0, p->stream>>>(
            p->view(), p->n_pairs, p->d_pi.p, p->d_pj.p, p->d_pw.p, p->d_pl.p, (float)contact_range, p->d_meta.p);
        CU(cudaGetLastError());
        kern<<<(int)grid, AD_THREADS, 0, p->stream>>>(p->view(), p->n_pairs, p->d_meta.p, p->d_dist.p, p->d_prob.p, p->d_valid.p);
    } else {
        const size_t smem = (size_t)p->max_copies * p->max_copies * p->n_struct * sizeof(float);
        if ((long long)smem + 4096 > (long long)p->smem_optin) {            // + the kernel's static shared memory
            char msg[256];
            snprintf(msg, sizeof msg, "igmk_pop_actdist: %d copy pairs x %d structures need %zu B of shared memory per CTA (limit %d)",
                     p->max_copies * p->max_copies, p->n_struct, smem, p->smem_optin);
            return fail(msg);
        }
        // fewer than three 256-thread CTAs per SM: CTAs of 1024 threads keep more loads in flight (measured at 8000
        // structures: 23.1 -> 9.2 ms per 101 k pairs; at 4000, where three narrow CTAs fit, the narrow form wins)
        bool wide = smem * 3 > (size_t)p->smem_optin;
        if (const char* e = getenv("IGMK_ACTDIST_NT")) wide = atoi(e) >= 1024;      // tuning aid
        auto kern = wide ? k_actdist<1024> : k_actdist<AD_THREADS>;
        const int nt = wide ? 1024 : AD_THREADS;
        CU(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        int per_sm = 1;
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nt, smem));
        if (per_sm < 1) per_sm = 1;
        const int64_t grid = std::min<int64_t>(p->n_pairs, (int64_t)p->sm_count * per_sm);
        CU(cudaEventRecord(p->ev0, p->stream));
        kern<<<(int)grid, nt, smem, p->stream>>>(p->view(), p->n_pairs, p->d_pi.p, p->d_pj.p, p->d_pw.p, p->d_pl.p,
                                                 (float)contact_range, p->d_dist.p, p->d_prob.p, p->d_valid.p);
    }
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    CU(cudaEventSynchronize(p->ev1));
    CU(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    p->actdist_done = true;
    return 0;
}

extern "C" int igmk_pop_actdist_collect(igmk_pop_t* p, int64_t max_rows, int32_t* row, int32_t* col, float* dist,
                                        float* prob, int64_t* n_rows_out)
{
    if (!p || !n_rows_out) return fail("igmk_pop_actdist_collect: NULL argument");
    *n_rows_out = 0;
    if (!p->actdist_done) return fail("igmk_pop_actdist_collect: no launch to collect");
    const int64_t n_pairs = p->n_pairs;
    if (n_pairs == 0) return 0;
    if (!row || !col || !dist || !prob) return fail("igmk_pop_actdist_collect: NULL argument");
    CU(cudaSetDevice(p->device));
    std::vector<float> hd(n_pairs), hp(n_pairs);
    std::vector<uint8_t> hv(n_pairs);
    CU(cudaMemcpyAsync(hd.data(), p->d_dist.p, n_pairs * sizeof(float), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaMemcpyAsync(hp.data(), p->d_prob.p, n_pairs * sizeof(float), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaMemcpyAsync(hv.data(), p->d_valid.p, n_pairs * sizeof(uint8_t), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    // expand to rows in the reference's emission order (ActivationDistanceStep.py:450-456)
    int64_t nr = 0;
    for (int64_t k = 0; k < n_pairs; k++) {
        if (!hv[k]) continue;
        const int hi = p->h_pi[k], hj = p->h_pj[k];
        const int io = p->copy_off[hi], jo = p->copy_off[hj];
        const int ni = p->copy_off[hi + 1] - io, nj = p->copy_off[hj + 1] - jo;
        const bool intra = p->chrom[hi] == p->chrom[hj];
        const int ne = intra ? std::min(ni, nj) : ni * nj;
        if (nr + ne > max_rows) return fail("igmk_pop_actdist_collect: output buffers too small");
        for (int c = 0; c < ne; c++) {
            row[nr] = p->copy_beads[io + (intra ? c : c / nj)];
            col[nr] = p->copy_beads[jo + (intra ? c : c % nj)];
            dist[nr] = hd[k]; prob[nr] = hp[k];
            nr++;
        }
    }
    *n_rows_out = nr;
    return 0;
}

extern "C" int igmk_pop_actdist(igmk_pop_t* p, int64_t n_pairs, const int32_t* pair_i, const int32_t* pair_j,
                                const double* pwish, const double* plast, double contact_range,
                                int64_t max_rows, int32_t* row, int32_t* col, float* dist, float* prob,
                                int64_t* n_rows_out)
{
    if (int rc = igmk_pop_set_pairs(p, n_pairs, pair_i, pair_j, pwish, plast)) return rc;
    if (int rc = igmk_pop_actdist_launch(p, contact_range)) return rc;
    return igmk_pop_actdist_collect(p, max_rows, row, col, dist, prob, n_rows_out);
}

// ---- activation distances with the population sharded by structure (astep_shard.cuh) -----------------
// The pairs are staged with igmk_pop_set_pairs on this rank's LOCAL population; the caller owns the device buffers
// and all-reduces (sum) d_counts after _count and d_hist after every _hist over the ranks (NCCL).  Every call has
// finished on the device when it returns.
static int shard_ready(igmk_pop_t* p, const char* who, const void* a, const void* b)
{
    if (!p) return fail(std::string(who) + ": NULL handle");
    if (p->n_pairs > 0 && (!a || !b)) return fail(std::string(who) + ": NULL device buffer");
    return 0;
}

extern "C" int igmk_pop_shard_count(igmk_pop_t* p, double contact_range, void* d_counts)
{
    if (int rc = shard_ready(p, "igmk_pop_shard_count", d_counts, d_counts)) return rc;
    p->actdist_done = false;
    if (p->n_pairs == 0) return 0;
    if (!p->d_crd) return fail("igmk_pop_shard_count: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const int64_t grid = std::min<int64_t>(p->n_pairs, (int64_t)p->sm_count * 8);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_shard_count<<<(int)grid, AD_THREADS, 0, p->stream>>>(p->view(), p->n_pairs, p->d_pi.p, p->d_pj.p, (float)contact_range, (uint32_t*)d_counts);
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    CU(cudaEventSynchronize(p->ev1));
    CU(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    return 0;
}

extern "C" int igmk_pop_shard_rank(igmk_pop_t* p, int32_t n_struct_total, const void* d_counts_total, void* d_want, void* d_prefix)
{
    if (int rc = shard_ready(p, "igmk_pop_shard_rank", d_counts_total, d_want)) return rc;
    if (n_struct_total < p->n_struct) return fail("igmk_pop_shard_rank: the whole population cannot be smaller than this rank's share");
    if (p->n_pairs == 0) return 0;
    if (!d_prefix) return fail("igmk_pop_shard_rank: NULL device buffer");
    CU(cudaSetDevice(p->device));
    k_shard_rank<<<(int)std::min<int64_t>((p->n_pairs + 255) / 256, (int64_t)p->sm_count * 8), 256, 0, p->stream>>>(
        p->view(), p->n_pairs, n_struct_total, p->d_pi.p, p->d_pj.p, p->d_pw.p, p->d_pl.p, (const uint32_t*)d_counts_total,
        (uint32_t*)d_want, (uint32_t*)d_prefix, p->d_prob.p, p->d_valid.p);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(p->stream));
    return 0;
}

extern "C" int igmk_pop_shard_hist(igmk_pop_t* p, int32_t pass, const void* d_prefix, void* d_hist)
{
    if (int rc = shard_ready(p, "igmk_pop_shard_hist", d_prefix, d_hist)) return rc;
    if (pass < 0 || pass > 3) return fail("igmk_pop_shard_hist: pass must be 0..3");
    if (p->n_pairs == 0) return 0;
    if (!p->d_crd) return fail("igmk_pop_shard_hist: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const int64_t grid = std::min<int64_t>(p->n_pairs, (int64_t)p->sm_count * 8);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_shard_hist<<<(int)grid, AD_THREADS, 0, p->stream>>>(p->view(), p->n_pairs, p->d_pi.p, p->d_pj.p, p->d_valid.p,
                                                          (const uint32_t*)d_prefix, pass, (uint32_t*)d_hist);
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    CU(cudaEventSynchronize(p->ev1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, p->ev0, p->ev1));
    p->last_ms = pass == 0 ? ms : p->last_ms + ms;        // the four histogram passes add up
    return 0;
}

extern "C" int igmk_pop_shard_pick(igmk_pop_t* p, int32_t pass, const void* d_hist_total, void* d_want, void* d_prefix)
{
    if (int rc = shard_ready(p, "igmk_pop_shard_pick", d_hist_total, d_want)) return rc;
    if (pass < 0 || pass > 3) return fail("igmk_pop_shard_pick: pass must be 0..3");
    if (p->n_pairs == 0) { p->actdist_done = pass == 3; return 0; }
    if (!d_prefix) return fail("igmk_pop_shard_pick: NULL device buffer");
    CU(cudaSetDevice(p->device));
    k_shard_pick<<<(int)std::min<int64_t>((p->n_pairs + 7) / 8, (int64_t)p->sm_count * 8), 256, 0, p->stream>>>(
        p->n_pairs, p->d_valid.p, (const uint32_t*)d_hist_total, pass, (uint32_t*)d_want, (uint32_t*)d_prefix, p->d_dist.p);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(p->stream));
    if (pass == 3) p->actdist_done = true;                // igmk_pop_actdist_collect expands the pairs to rows
    return 0;
}

extern "C" int igmk_pop_contact_freq(igmk_pop_t* p, int64_t n_pairs, const int32_t* pair_i, const int32_t* pair_j,
                                     double contact_range, float* freq)
{
    if (!p) return fail("igmk_pop_contact_freq: NULL handle");
    if (n_pairs == 0) return 0;
    if (!freq) return fail("igmk_pop_contact_freq: NULL argument");
    const double zero = 0.0;
    (void)zero;
    std::vector<double> dummy((size_t)n_pairs, 0.0);
    if (int rc = igmk_pop_set_pairs(p, n_pairs, pair_i, pair_j, dummy.data(), dummy.data())) return rc;
    if (!p->d_crd) return fail("igmk_pop_contact_freq: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const int64_t grid = std::min<int64_t>((n_pairs + 7) / 8, (int64_t)p->sm_count * 32);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_contact_freq<<<(int)grid, 256, 0, p->stream>>>(p->view(), n_pairs, p->d_pi.p, p->d_pj.p, (float)contact_range, p->d_dist.p);
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    CU(cudaMemcpyAsync(freq, p->d_dist.p, n_pairs * sizeof(float), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    CU(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    p->n_pairs = 0; p->actdist_done = false;
    return 0;
}

// ---- DamID (lamina) A-step -----------------------------------------------------------------
static void damid_denominators(const igmk_pop* p, int64_t n, const int32_t* bead_of, double contact_range, int ellipsoid,
                               const double* nucleus_param, const std::vector<float>& radii, std::vector<double>& den)
{
    den.resize((size_t)n * 3);
    for (int64_t q = 0; q < n; q++) {
        const double r = (double)radii[bead_of[q]];
        for (int k = 0; k < 3; k++) {
            const double axis = nucleus_param[ellipsoid ? k : 0] * (1 - contact_range);
            den[3 * q + k] = pow(axis - r, 2.0);          // numpy scalar `**2` == libm pow
        }
    }
    (void)p;
}

extern "C" int igmk_pop_damid_actdist(igmk_pop_t* p, int64_t n_loci, const int32_t* loc, const double* pwish, const double* plast,
                                      double contact_range, int32_t ellipsoid, const double* nucleus_param,
                                      int64_t max_rows, int32_t* row_bead, float* dist, float* prob, int64_t* n_rows_out)
{
    if (!p || !n_rows_out) return fail("igmk_pop_damid_actdist: NULL argument");
    *n_rows_out = 0;
    if (n_loci == 0) return 0;
    if (!loc || !pwish || !plast || !nucleus_param || !row_bead || !dist || !prob) return fail("igmk_pop_damid_actdist: NULL argument");
    for (int64_t q = 0; q < n_loci; q++)
        if (loc[q] < 0 || loc[q] >= p->n_hap) return fail("igmk_pop_damid_actdist: locus index out of range");
    if (!p->d_crd) return fail("igmk_pop_damid_actdist: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const size_t smem = (size_t)p->max_copies * p->n_struct * sizeof(double);
    if ((long long)smem + 1024 > (long long)p->smem_optin) return fail("igmk_pop_damid_actdist: population too large for one CTA's shared memory");
    std::vector<int32_t> bead0((size_t)n_loci);
    for (int64_t q = 0; q < n_loci; q++) bead0[q] = p->copy_beads[p->copy_off[loc[q]]];
    std::vector<double> den;
    damid_denominators(p, n_loci, bead0.data(), contact_range, ellipsoid, nucleus_param, p->radii, den);
    CU(p->d_pi.reserve(n_loci)); CU(p->d_pw.reserve(n_loci)); CU(p->d_pl.reserve(n_loci));
    CU(p->d_dist.reserve(n_loci)); CU(p->d_prob.reserve(n_loci)); CU(p->d_den.reserve((size_t)n_loci * 3));
    CU(cudaMemcpyAsync(p->d_pi.p, loc, n_loci * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_pw.p, pwish, n_loci * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_pl.p, plast, n_loci * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_den.p, den.data(), den.size() * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    CU(cudaFuncSetAttribute(k_damid_actdist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t grid = std::min<int64_t>(n_loci, (int64_t)p->sm_count * 8);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_damid_actdist<<<(int)grid, AD_THREADS, smem, p->stream>>>(p->view(), n_loci, p->d_pi.p, p->d_pw.p, p->d_pl.p, p->d_den.p,
                                                                 ellipsoid ? 1 : 0, p->d_dist.p, p->d_prob.p);
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    std::vector<float> hd((size_t)n_loci), hp((size_t)n_loci);
    CU(cudaMemcpyAsync(hd.data(), p->d_dist.p, n_loci * sizeof(float), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaMemcpyAsync(hp.data(), p->d_prob.p, n_loci * sizeof(float), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    CU(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    p->n_pairs = 0; p->actdist_done = false;
    // one row per copy of the locus, always (DamidActivationDistanceStep.py:356)
    int64_t nr = 0;
    for (int64_t q = 0; q < n_loci; q++) {
        const int o0 = p->copy_off[loc[q]], nc = p->copy_off[loc[q] + 1] - o0;
        if (nr + nc > max_rows) return fail("igmk_pop_damid_actdist: output buffers too small");
        for (int c = 0; c < nc; c++) { row_bead[nr] = p->copy_beads[o0 + c]; dist[nr] = hd[q]; prob[nr] = hp[q]; nr++; }
    }
    *n_rows_out = nr;
    return 0;
}

extern "C" int igmk_pop_damid_assign(igmk_pop_t* p, int64_t n_rows, const int32_t* bead, const float* dist,
                                     double contact_range, const double* semiaxes, uint32_t* bits)
{
    if (!p) return fail("igmk_pop_damid_assign: NULL handle");
    if (n_rows == 0) return 0;
    if (!bead || !dist || !semiaxes || !bits) return fail("igmk_pop_damid_assign: NULL argument");
    for (int64_t r = 0; r < n_rows; r++)
        if (bead[r] < 0 || bead[r] >= p->n_bead) return fail("igmk_pop_damid_assign: bead index out of range");
    if (!p->d_crd) return fail("igmk_pop_damid_assign: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    std::vector<double> den;
    damid_denominators(p, n_rows, bead, contact_range, 1, semiaxes, p->radii, den);
    const int nw = (p->n_struct + 31) / 32;
    CU(p->d_row.reserve(n_rows)); CU(p->d_rdist.reserve(n_rows)); CU(p->d_den.reserve((size_t)n_rows * 3));
    CU(p->d_bits.reserve((size_t)n_rows * nw));
    CU(cudaMemcpyAsync(p->d_row.p, bead, n_rows * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_rdist.p, dist, n_rows * sizeof(float), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_den.p, den.data(), den.size() * sizeof(double), cudaMemcpyHostToDevice, p->stream));
    const int64_t grid = std::min<int64_t>((n_rows + 31) / 32, (int64_t)p->sm_count * 2);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_damid_assign<<<(int)grid, 1024, 0, p->stream>>>(p->view(), n_rows, p->d_row.p, p->d_rdist.p, p->d_den.p, p->d_bits.p);
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    CU(cudaMemcpyAsync(bits, p->d_bits.p, (size_t)n_rows * nw * sizeof(uint32_t), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    CU(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    p->n_rows = 0; p->assign_done = false;
    return 0;
}

// ---- FISH A-step ---------------------------------------------------------------------------
extern "C" int igmk_pop_fish_ranks(igmk_pop_t* p, int64_t n_items, const int32_t* hap_i, const int32_t* hap_j,
                                   float* min_dist, float* max_dist, int32_t* rank_min, int32_t* rank_max)
{
    if (!p) return fail("igmk_pop_fish_ranks: NULL handle");
    if (n_items == 0) return 0;
    if (!hap_i || !hap_j || !min_dist || !max_dist || !rank_min || !rank_max) return fail("igmk_pop_fish_ranks: NULL argument");
    for (int64_t q = 0; q < n_items; q++) {
        if (hap_i[q] < 0 || hap_i[q] >= p->n_hap || hap_j[q] >= p->n_hap) return fail("igmk_pop_fish_ranks: locus index out of range");
        // the reference indexes copies [0] and [1] (FishAssignmentStep.py:41-44, :63-64): diploid loci only
        if (p->copy_off[hap_i[q] + 1] - p->copy_off[hap_i[q]] < 2 ||
            (hap_j[q] >= 0 && p->copy_off[hap_j[q] + 1] - p->copy_off[hap_j[q]] < 2))
            return fail("igmk_pop_fish_ranks: locus has fewer than two copies (IndexError in the reference)");
    }
    if (!p->d_crd) return fail("igmk_pop_fish_ranks: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const size_t S = (size_t)p->n_struct;
    const size_t smem = 2 * S * sizeof(float);
    if ((long long)smem + 1024 > (long long)p->smem_optin) return fail("igmk_pop_fish_ranks: population too large for one CTA's shared memory");
    CU(p->d_pi.reserve(n_items)); CU(p->d_pj.reserve(n_items));
    DevBuf<float> d_lo, d_hi; DevBuf<int32_t> d_ra, d_rb;
    CU(d_lo.reserve(n_items * S)); CU(d_hi.reserve(n_items * S)); CU(d_ra.reserve(n_items * S)); CU(d_rb.reserve(n_items * S));
    CU(cudaMemcpyAsync(p->d_pi.p, hap_i, n_items * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_pj.p, hap_j, n_items * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaFuncSetAttribute(k_fish_ranks, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t grid = std::min<int64_t>(n_items, (int64_t)p->sm_count * 8);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_fish_ranks<<<(int)grid, AD_THREADS, smem, p->stream>>>(p->view(), n_items, p->d_pi.p, p->d_pj.p, d_lo.p, d_hi.p, d_ra.p, d_rb.p);
    cudaError_t le = cudaGetLastError();
    if (le == cudaSuccess) le = cudaEventRecord(p->ev1, p->stream);
    if (le == cudaSuccess) le = cudaMemcpyAsync(min_dist, d_lo.p, n_items * S * sizeof(float), cudaMemcpyDeviceToHost, p->stream);
    if (le == cudaSuccess) le = cudaMemcpyAsync(max_dist, d_hi.p, n_items * S * sizeof(float), cudaMemcpyDeviceToHost, p->stream);
    if (le == cudaSuccess) le = cudaMemcpyAsync(rank_min, d_ra.p, n_items * S * sizeof(int32_t), cudaMemcpyDeviceToHost, p->stream);
    if (le == cudaSuccess) le = cudaMemcpyAsync(rank_max, d_rb.p, n_items * S * sizeof(int32_t), cudaMemcpyDeviceToHost, p->stream);
    if (le == cudaSuccess) le = cudaStreamSynchronize(p->stream);
    if (le == cudaSuccess) le = cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1);
    d_lo.release(); d_hi.release(); d_ra.release(); d_rb.release();
    CU(le);
    p->n_pairs = 0; p->actdist_done = false;
    return 0;
}

// ---- SPRITE A-step ---------------------------------------------------------------------------
extern "C" int igmk_pop_sprite_rg2(igmk_pop_t* p, int32_t n_clusters, const int32_t* cl_off, const int32_t* rg_off,
                                   const int32_t* beads, float* rg2, int8_t* copy_idx, int32_t* min_struct)
{
    if (!p) return fail("igmk_pop_sprite_rg2: NULL handle");
    if (n_clusters == 0) return 0;
    if (!cl_off || !rg_off || !beads || !rg2 || !copy_idx || !min_struct) return fail("igmk_pop_sprite_rg2: NULL argument");
    if (cl_off[0] != 0 || rg_off[0] != 0) return fail("igmk_pop_sprite_rg2: offsets must start at 0");
    const int32_t n_regions = cl_off[n_clusters];
    for (int c = 0; c < n_clusters; c++) {
        const int nr = cl_off[c + 1] - cl_off[c];
        if (nr < 1 || nr > SPRITE_MAX_REGIONS) return fail("igmk_pop_sprite_rg2: a cluster needs 1..8 regions");
        long long comb = 1;
        for (int i = cl_off[c]; i < cl_off[c + 1]; i++) {
            const int k = rg_off[i + 1] - rg_off[i];
            if (k < 1 || k > 127) return fail("igmk_pop_sprite_rg2: a region needs 1..127 alternative beads");
            comb *= k;
        }
        if (comb > 65536) return fail("igmk_pop_sprite_rg2: more than 65536 copy combinations in one cluster");
    }
    const int32_t n_beads = rg_off[n_regions];
    for (int i = 0; i < n_beads; i++)
        if (beads[i] < 0 || beads[i] >= p->n_bead) return fail("igmk_pop_sprite_rg2: bead index out of range");
    if (!p->d_crd) return fail("igmk_pop_sprite_rg2: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const size_t S = (size_t)p->n_struct;
    DevBuf<int32_t> d_cl, d_rg, d_b, d_ms; DevBuf<float> d_v; DevBuf<int8_t> d_ci;
    cudaError_t e = d_cl.reserve(n_clusters + 1);
    if (e == cudaSuccess) e = d_rg.reserve(n_regions + 1);
    if (e == cudaSuccess) e = d_b.reserve(std::max(n_beads, 1));
    if (e == cudaSuccess) e = d_ms.reserve(n_clusters);
    if (e == cudaSuccess) e = d_v.reserve((size_t)n_clusters * S);
    if (e == cudaSuccess) e = d_ci.reserve((size_t)n_regions * S);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_cl.p, cl_off, (n_clusters + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rg.p, rg_off, (n_regions + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_b.p, beads, n_beads * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaEventRecord(p->ev0, p->stream);
    if (e == cudaSuccess) {
        const int grid = std::min<int>(n_clusters, p->sm_count * 8);
        k_sprite_rg2<<<grid, 256, 0, p->stream>>>(p->view(), n_clusters, d_cl.p, d_rg.p, d_b.p, d_v.p, d_ci.p, d_ms.p);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaEventRecord(p->ev1, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rg2, d_v.p, (size_t)n_clusters * S * sizeof(float), cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(copy_idx, d_ci.p, (size_t)n_regions * S, cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(min_struct, d_ms.p, n_clusters * sizeof(int32_t), cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1);
    d_cl.release(); d_rg.release(); d_b.release(); d_ms.release(); d_v.release(); d_ci.release();
    CU(e);
    return 0;
}

extern "C" int igmk_pop_sprite_select(igmk_pop_t* p, int32_t n_clusters, const int32_t* cl_off, const int32_t* seg_off,
                                      const int32_t* beads, const int32_t* seg_group, const int32_t* rep, int32_t keep,
                                      float* best_rg2, int32_t* best_struct, int32_t* best_beads)
{
    if (!p) return fail("igmk_pop_sprite_select: NULL handle");
    if (n_clusters == 0) return 0;
    if (!cl_off || !seg_off || !beads || !seg_group || !rep || !best_rg2 || !best_struct || !best_beads)
        return fail("igmk_pop_sprite_select: NULL argument");
    if (cl_off[0] != 0 || seg_off[0] != 0) return fail("igmk_pop_sprite_select: offsets must start at 0");
    if (keep < 1 || keep >= p->n_struct) return fail("igmk_pop_sprite_select: keep_best must be 1 .. n_struct - 1 (np.argpartition, SpriteAssignmentStep.py:148)");
    if (p->n_struct > SPRITE_SORT_MAX) return fail("igmk_pop_sprite_select: more than 4096 structures");
    if (!p->d_crd) return fail("igmk_pop_sprite_select: no coordinates uploaded or attached");
    const int32_t n_seg = cl_off[n_clusters];
    std::vector<int32_t> ngroups((size_t)n_clusters);
    for (int c = 0; c < n_clusters; c++) {
        if (cl_off[c + 1] <= cl_off[c]) return fail("igmk_pop_sprite_select: empty cluster");
        int ng = 0;
        for (int i = cl_off[c]; i < cl_off[c + 1]; i++) {
            const int k = seg_off[i + 1] - seg_off[i];
            if (k < 1 || k > 15) return fail("igmk_pop_sprite_select: a segment needs 1..15 copies");
            if (seg_group[i] < 0 || seg_group[i] >= SPRITE_MAX_GROUPS) return fail("igmk_pop_sprite_select: more than 8 chromosomes in one cluster");
            ng = std::max(ng, seg_group[i] + 1);
        }
        ngroups[c] = ng;
        long long comb = 1;
        for (int g = 0; g < ng; g++) {
            const int r = rep[c * SPRITE_MAX_GROUPS + g];
            if (r < 0 || r >= cl_off[c + 1] - cl_off[c] || seg_group[cl_off[c] + r] != g)
                return fail("igmk_pop_sprite_select: representative segment does not belong to its group");
            comb *= seg_off[cl_off[c] + r + 1] - seg_off[cl_off[c] + r];
        }
        if (comb > 65536) return fail("igmk_pop_sprite_select: more than 65536 copy combinations in one cluster");
    }
    const int32_t n_beads = seg_off[n_seg];
    for (int i = 0; i < n_beads; i++)
        if (beads[i] < 0 || beads[i] >= p->n_bead) return fail("igmk_pop_sprite_select: bead index out of range");
    CU(cudaSetDevice(p->device));
    DevBuf<int32_t> d_cl, d_so, d_b, d_sg, d_ng, d_rep, d_bs, d_bb; DevBuf<float> d_v;
    cudaError_t e = d_cl.reserve(n_clusters + 1);
    if (e == cudaSuccess) e = d_so.reserve(n_seg + 1);
    if (e == cudaSuccess) e = d_b.reserve(std::max(n_beads, 1));
    if (e == cudaSuccess) e = d_sg.reserve(n_seg);
    if (e == cudaSuccess) e = d_ng.reserve(n_clusters);
    if (e == cudaSuccess) e = d_rep.reserve((size_t)n_clusters * SPRITE_MAX_GROUPS);
    if (e == cudaSuccess) e = d_v.reserve((size_t)n_clusters * keep);
    if (e == cudaSuccess) e = d_bs.reserve((size_t)n_clusters * keep);
    if (e == cudaSuccess) e = d_bb.reserve((size_t)n_seg * keep);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_cl.p, cl_off, (n_clusters + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_so.p, seg_off, (n_seg + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_b.p, beads, n_beads * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_sg.p, seg_group, n_seg * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_ng.p, ngroups.data(), n_clusters * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rep.p, rep, (size_t)n_clusters * SPRITE_MAX_GROUPS * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream);
    int npow = 1;
    while (npow < p->n_struct) npow <<= 1;
    const size_t smem = (size_t)npow * 8 + (size_t)p->n_struct * 4;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_sprite_select, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) e = cudaEventRecord(p->ev0, p->stream);
    if (e == cudaSuccess) {
        k_sprite_select<<<std::min<int>(n_clusters, p->sm_count * 4), 256, smem, p->stream>>>(
            p->view(), n_clusters, d_cl.p, d_so.p, d_b.p, d_sg.p, d_ng.p, d_rep.p, keep, d_v.p, d_bs.p, d_bb.p);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaEventRecord(p->ev1, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_rg2, d_v.p, (size_t)n_clusters * keep * sizeof(float), cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_struct, d_bs.p, (size_t)n_clusters * keep * sizeof(int32_t), cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_beads, d_bb.p, (size_t)n_seg * keep * sizeof(int32_t), cudaMemcpyDeviceToHost, p->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(p->stream);
    if (e == cudaSuccess) e = cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1);
    d_cl.release(); d_so.release(); d_b.release(); d_sg.release(); d_ng.release(); d_rep.release(); d_v.release(); d_bs.release(); d_bb.release();
    CU(e);
    return 0;
}

extern "C" int igmk_pop_set_rows(igmk_pop_t* p, int64_t n_rows, const int32_t* row, const int32_t* col, const float* dist)
{
    if (!p) return fail("igmk_pop_set_rows: NULL handle");
    if (n_rows < 0) return fail("igmk_pop_set_rows: negative row count");
    p->n_rows = 0; p->assign_done = false; p->bounds_rows = -1;
    if (n_rows == 0) return 0;
    if (!row || !col || !dist) return fail("igmk_pop_set_rows: NULL argument");
    for (int64_t r = 0; r < n_rows; r++)
        if (row[r] < 0 || row[r] >= p->n_bead || col[r] < 0 || col[r] >= p->n_bead)
            return fail("igmk_pop_set_rows: bead index out of range");
    CU(cudaSetDevice(p->device));
    const int nw = (p->n_struct + 31) / 32;
    CU(p->d_row.reserve(n_rows)); CU(p->d_col.reserve(n_rows)); CU(p->d_rdist.reserve(n_rows));
    CU(p->d_bits.reserve((size_t)n_rows * nw));
    CU(cudaMemcpyAsync(p->d_row.p, row, n_rows * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_col.p, col, n_rows * sizeof(int32_t), cudaMemcpyHostToDevice, p->stream));
    CU(cudaMemcpyAsync(p->d_rdist.p, dist, n_rows * sizeof(float), cudaMemcpyHostToDevice, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    p->n_rows = n_rows;
    return 0;
}

extern "C" int igmk_pop_assign_launch(igmk_pop_t* p)
{
    if (!p) return fail("igmk_pop_assign_launch: NULL handle");
    p->last_ms = 0.f;
    if (p->n_rows == 0) { p->assign_done = true; return 0; }
    if (!p->d_crd) return fail("igmk_pop_assign_launch: no coordinates uploaded or attached");
    CU(cudaSetDevice(p->device));
    const int64_t grid = std::min<int64_t>((p->n_rows + 31) / 32, (int64_t)p->sm_count * 2);
    CU(cudaEventRecord(p->ev0, p->stream));
    k_assign<<<(int)grid, 1024, 0, p->stream>>>(p->view(), p->n_rows, p->d_row.p, p->d_col.p, p->d_rdist.p, p->d_bits.p);
    CU(cudaGetLastError());
    CU(cudaEventRecord(p->ev1, p->stream));
    CU(cudaEventSynchronize(p->ev1));
    CU(cudaEventElapsedTime(&p->last_ms, p->ev0, p->ev1));
    p->assign_done = true;
    return 0;
}

extern "C" int igmk_pop_assign_collect(igmk_pop_t* p, uint32_t* bits)
{
    if (!p) return fail("igmk_pop_assign_collect: NULL handle");
    if (!p->assign_done) return fail("igmk_pop_assign_collect: no launch to collect");
    if (p->n_rows == 0) return 0;
    if (!bits) return fail("igmk_pop_assign_collect: NULL argument");
    CU(cudaSetDevice(p->device));
    const int nw = (p->n_struct + 31) / 32;
    CU(cudaMemcpyAsync(bits, p->d_bits.p, (size_t)p->n_rows * nw * sizeof(uint32_t), cudaMemcpyDeviceToHost, p->stream));
    CU(cudaStreamSynchronize(p->stream));
    return 0;
}

extern "C" int igmk_pop_assign(igmk_pop_t* p, int64_t n_rows, const int32_t* row, const int32_t* col,
                               const float* dist, uint32_t* bits)
{
    if (int rc = igmk_pop_set_rows(p, n_rows, row, col, dist)) return rc;
    if (int rc = igmk_pop_assign_launch(p)) return rc;
    return igmk_pop_assign_collect(p, bits);
}

// =========================================================================================
// Assignment -> M-step batch without leaving the device (assemble.cuh)

extern "C" int igmk_pop_set_row_bounds(igmk_pop_t* p, int64_t n_rows, const double* row_r0, double row_k)
{
    if (!p) return fail("igmk_pop_set_row_bounds: NULL handle");
    if (n_rows != p->n_rows) return fail("igmk_pop_set_row_bounds: one bound per row staged with igmk_pop_set_rows is needed");
    if (!(row_k >= 0.0)) return fail("igmk_pop_set_row_bounds: the spring constant must be >= 0");
    p->bounds_rows = -1; p->key_r0.clear(); p->row_k = row_k;
    if (n_rows == 0) { p->bounds_rows = 0; return 0; }
    if (!row_r0) return fail("igmk_pop_set_row_bounds: NULL argument");
    std::vector<int32_t> key((size_t)n_rows);
    std::map<double, int32_t> seen;
    int32_t last = -1;
    for (int64_t r = 0; r < n_rows; r++) {
        if (last >= 0 && row_r0[r] == row_r0[r - 1]) { key[r] = last; continue; }     // runs of equal bounds skip the map
        if (!(row_r0[r] >= 0.0)) return fail("igmk_pop_set_row_bounds: negative bound distance");
        auto it = seen.find(row_r0[r]);
        if (it == seen.end()) { last = (int32_t)p->key_r0.size(); seen[row_r0[r]] = last; p->key_r0.push_back(row_r0[r]); }
        else last = it->second;
        key[r] = last;
    }
    CU(cudaSetDevice(p->device));
    CU(p->d_rkey.reserve((size_t)n_rows));
    CU(copy_on(p->stream, p->d_rkey.p, key.data(), (size_t)n_rows * sizeof(int32_t), cudaMemcpyHostToDevice));
    p->bounds_rows = n_rows;
    return 0;
}

// neighbour tables and work areas of a device-assembled batch (what upload() allocates after its copies)
static int alloc_run_buffers(igmk_handle* h)
{
    const size_t na = h->xyzr.size(), ns = h->desc.size();
    CU(cudaMalloc((void**)&h->d_nl, std::max<size_t>(na, 1) * h->nl_width * sizeof(uint16_t)));
    CU(cudaMalloc((void**)&h->d_md2_nl, ns * (size_t)h->md2_K * MD2S_T * (size_t)h->md2_nlw * sizeof(uint16_t)));
    return 0;
}

// after a neighbour-table overflow: wider tables, submitted coordinates back in place
static int upload_dev_rebuild(igmk_handle* h)
{
    cudaFree(h->d_nl); cudaFree(h->d_md2_nl); h->d_nl = nullptr; h->d_md2_nl = nullptr;
    if (int rc = alloc_run_buffers(h)) return rc;
    CU(copy_on(h->stream, h->d_xyzr, h->d_xyzr0, h->xyzr.size() * sizeof(float4), cudaMemcpyDeviceToDevice));
    h->dirty = false;
    return 0;
}

static int submit_population_dev_impl(igmk_handle_t* h, const igmk_system_t* tmpl, igmk_pop_t* pop, int32_t s_begin,
                                      int32_t n_struct, const int64_t* seeds, int32_t* first_slot_out);

extern "C" int igmk_submit_population_dev(igmk_handle_t* h, const igmk_system_t* tmpl, igmk_pop_t* pop, int32_t s_begin,
                                          int32_t n_struct, const int64_t* seeds, int32_t* first_slot_out)
{
    if (!h || !tmpl || !pop) return fail("igmk_submit_population_dev: NULL argument");
    if (!h->desc.empty()) return fail("igmk_submit_population_dev: the batch must be empty (igmk_clear)");
    const int rc = submit_population_dev_impl(h, tmpl, pop, s_begin, n_struct, seeds, first_slot_out);
    if (rc != 0) {                         // leave no half-assembled batch behind (the message of the failure is kept)
        const std::string msg = g_err;
        igmk_clear(h);
        g_err = msg;
    }
    return rc;
}

static int submit_population_dev_impl(igmk_handle_t* h, const igmk_system_t* tmpl, igmk_pop_t* pop, int32_t s_begin,
                                      int32_t n_struct, const int64_t* seeds, int32_t* first_slot_out)
{
    if (h->device < 0 || h->device != pop->device) return fail("igmk_submit_population_dev: batch and population must live on the same GPU");
    if (n_struct <= 0 || s_begin < 0 || s_begin + n_struct > pop->n_struct) return fail("igmk_submit_population_dev: bad structure range");
    if (!pop->d_crd) return fail("igmk_submit_population_dev: the population has no coordinates");
    if (pop->n_rows > 0 && !pop->assign_done) return fail("igmk_submit_population_dev: no assignment on the device (igmk_pop_assign_launch)");
    if (pop->bounds_rows != pop->n_rows) return fail("igmk_submit_population_dev: no bound distances for the staged rows (igmk_pop_set_row_bounds)");
    const int n = tmpl->n_atoms, n_bead = pop->n_bead;
    if (n_bead > n) return fail("igmk_submit_population_dev: the template has fewer atoms than the population has beads");
    // what this path does not cover is left to igmk_submit_population (status 2: the caller falls back)
    if (n > ASM_MAX_ATOMS || md2_pick_K(n) == 0 || getenv("IGMK_FORCE_BIG") || getenv("IGMK_FORCE_CLUSTER")) {
        fail("igmk_submit_population_dev: structures of this size are assembled on the host");
        return 2;
    }
    if (tmpl->n_env < 0 || tmpl->n_env > IGMK_MAX_ENVELOPES) return fail("igmk_submit_population_dev: too many envelopes");
    {
        const size_t need = std::max(ctx_smem_bytes<float>(n, true), ctx_smem_bytes<double>(n, false));
        if ((long long)need > (long long)h->smem_optin) return fail("igmk_submit_population_dev: structure too large for one CTA");
    }
    CU(cudaSetDevice(h->device));
    const bool trace_sp = getenv("IGMK_TRACE") != nullptr;
    auto tnow = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double tsp0 = tnow();
    // ---- the template: atoms, envelope masks, bonds, and the table of distinct (r0, k, style) over rows and template ----
    std::vector<float4> t_xyzr((size_t)n);
    std::vector<uint8_t> t_kind(tmpl->kind, tmpl->kind + n);
    std::vector<uint32_t> t_mask((size_t)n, 0u);
    float maxrad = 0.f; int nonfixed = 0;
    for (int i = 0; i < n; i++) {
        if (tmpl->kind[i] > 2) return fail("igmk_submit_population_dev: unknown particle kind");
        t_xyzr[i] = make_float4(tmpl->xyz[3 * i], tmpl->xyz[3 * i + 1], tmpl->xyz[3 * i + 2], tmpl->radius[i]);
        if (tmpl->kind[i] == IGMK_KIND_BEAD) maxrad = std::max(maxrad, tmpl->radius[i]);
        if (tmpl->kind[i] != IGMK_KIND_STATIC) nonfixed++;
    }
    for (int e = 0; e < tmpl->n_env; e++)
        for (int q = tmpl->env_off[e]; q < tmpl->env_off[e + 1]; q++) {
            const int i = tmpl->env_members[q];
            if (i < 0 || i >= n) return fail("igmk_submit_population_dev: envelope member out of range");
            t_mask[i] |= 1u << e;
        }
    std::vector<double> key_r0(pop->key_r0), key_k(pop->key_r0.size(), pop->row_k);
    std::vector<int32_t> key_style(pop->key_r0.size(), IGMK_BOND_UPPER);
    std::map<std::tuple<double, double, int>, int32_t> keys;
    for (size_t g = 0; g < key_r0.size(); g++) keys[std::make_tuple(key_r0[g], key_k[g], (int)IGMK_BOND_UPPER)] = (int32_t)g;
    const int nt = tmpl->n_bonds;
    std::vector<int32_t> t_i((size_t)nt), t_j((size_t)nt), t_key((size_t)nt);
    for (int b = 0; b < nt; b++) {
        const int i = tmpl->bond_i[b], j = tmpl->bond_j[b];
        if (i < 0 || i >= n || j < 0 || j >= n) return fail("igmk_submit_population_dev: bond index out of range");
        if (tmpl->bond_style[b] > 1) return fail("igmk_submit_population_dev: unknown bond style");
        if (!(tmpl->bond_k[b] >= 0.0)) return fail("igmk_submit_population_dev: bond spring constants must be >= 0");
        if (!(tmpl->bond_r0[b] >= 0.0)) return fail("igmk_submit_population_dev: negative bound distance in the template");
        const auto key = std::make_tuple(tmpl->bond_r0[b], tmpl->bond_k[b], (int)tmpl->bond_style[b]);
        auto it = keys.find(key);
        if (it == keys.end()) {
            it = keys.emplace(key, (int32_t)key_r0.size()).first;
            key_r0.push_back(tmpl->bond_r0[b]); key_k.push_back(tmpl->bond_k[b]); key_style.push_back((int32_t)tmpl->bond_style[b]);
        }
        t_i[b] = i; t_j[b] = j; t_key[b] = it->second;
    }
    const int G = (int)key_r0.size();
    if (G > ASM_MAX_KEYS) {
        fail("igmk_submit_population_dev: more distinct bound distances than the device path handles; assemble on the host");
        return 2;
    }
    // one device blob for all of it (offsets 16-byte aligned)
    auto al = [](size_t v) { return (v + 15) / 16 * 16; };
    size_t o_xyzr = 0, o_kind = al(o_xyzr + (size_t)n * sizeof(float4)), o_mask = al(o_kind + (size_t)n), o_ti = al(o_mask + (size_t)n * 4),
           o_tj = al(o_ti + (size_t)nt * 4), o_tk = al(o_tj + (size_t)nt * 4), o_kr = al(o_tk + (size_t)nt * 4), o_kk = al(o_kr + (size_t)G * 8),
           o_ks = al(o_kk + (size_t)G * 8), total = al(o_ks + (size_t)G * 4);
    std::vector<unsigned char> blob(total, 0);
    memcpy(blob.data() + o_xyzr, t_xyzr.data(), (size_t)n * sizeof(float4));
    memcpy(blob.data() + o_kind, t_kind.data(), (size_t)n);
    memcpy(blob.data() + o_mask, t_mask.data(), (size_t)n * 4);
    if (nt) { memcpy(blob.data() + o_ti, t_i.data(), (size_t)nt * 4); memcpy(blob.data() + o_tj, t_j.data(), (size_t)nt * 4); memcpy(blob.data() + o_tk, t_key.data(), (size_t)nt * 4); }
    if (G) { memcpy(blob.data() + o_kr, key_r0.data(), (size_t)G * 8); memcpy(blob.data() + o_kk, key_k.data(), (size_t)G * 8); memcpy(blob.data() + o_ks, key_style.data(), (size_t)G * 4); }
    h->free_device();
    CU(cudaMalloc((void**)&h->d_asm_blob, total));
    CU(cudaMemcpyAsync(h->d_asm_blob, blob.data(), total, cudaMemcpyHostToDevice, h->stream));
    CU(cudaMalloc((void**)&h->d_asm_cnt, (size_t)3 * n_struct * sizeof(int32_t)));
    AsmIn I;
    memset(&I, 0, sizeof I);
    I.n_atoms = n; I.n_bead = n_bead; I.n_tbonds = nt; I.n_keys = G;
    I.t_xyzr = (const float4*)(h->d_asm_blob + o_xyzr); I.t_kind = h->d_asm_blob + o_kind; I.t_envmask = (const uint32_t*)(h->d_asm_blob + o_mask);
    I.t_i = (const int32_t*)(h->d_asm_blob + o_ti); I.t_j = (const int32_t*)(h->d_asm_blob + o_tj); I.t_key = (const int32_t*)(h->d_asm_blob + o_tk);
    I.key_r0 = (const double*)(h->d_asm_blob + o_kr); I.key_k = (const double*)(h->d_asm_blob + o_kk); I.key_style = (const int32_t*)(h->d_asm_blob + o_ks);
    h->d_asm_key_r0 = I.key_r0; h->d_asm_key_k = I.key_k; h->d_asm_key_style = I.key_style;
    I.crd = pop->d_crd; I.n_struct_total = pop->n_struct; I.s_begin = s_begin;
    I.n_rows = pop->n_rows; I.row = pop->d_row.p; I.col = pop->d_col.p; I.row_key = pop->d_rkey.p;
    I.bits = pop->d_bits.p; I.bits_words = (pop->n_struct + 31) / 32;
    // ---- sizes ----
    k_asm_count<<<n_struct, ASM_T, 0, h->stream>>>(I, h->d_asm_cnt, h->d_asm_cnt + n_struct);
    CU(cudaGetLastError());
    std::vector<int32_t> cnt((size_t)2 * n_struct);
    CU(copy_on(h->stream, cnt.data(), h->d_asm_cnt, cnt.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    h->desc.resize((size_t)n_struct); h->struct_id.resize((size_t)n_struct); h->sub_bond_start.resize((size_t)n_struct);
    h->dev_nb.assign(cnt.begin(), cnt.begin() + n_struct);
    int64_t bonds = 0; int32_t btypes = 0;
    for (int sl = 0; sl < n_struct; sl++) {
        StructDesc d;
        memset(&d, 0, sizeof d);
        d.atom_off = sl * n; d.n_atoms = n; d.n_nonfixed = nonfixed;
        d.bond_ptr_off = sl * (n + 1); d.bond_item_off = 2 * bonds; d.btype_off = btypes;
        d.n_env = tmpl->n_env;
        for (int e = 0; e < tmpl->n_env; e++) for (int k = 0; k < 4; k++) d.env[e][k] = tmpl->env_abck[4 * e + k];
        d.evfactor = tmpl->evfactor;
        const int64_t sid = tmpl->struct_id + s_begin + sl;
        d.seed = seeds ? seeds[sl] : (int64_t)(((__int128)std::max<int64_t>(tmpl->seed, 1) * sid) % 9190037) + 1;
        d.maxrad = maxrad;
        h->desc[sl] = d; h->struct_id[sl] = sid; h->sub_bond_start[sl] = bonds;
        bonds += cnt[sl]; btypes += std::max(cnt[(size_t)n_struct + sl], 2);
    }
    h->dev_bonds = bonds; h->max_atoms = n;
    if ((size_t)n_struct * (size_t)(n + 1) > (size_t)INT32_MAX) return fail("igmk_submit_population_dev: batch too large");
    const size_t na = (size_t)n_struct * n;
    h->md2_K = md2_pick_K(n); h->md_big = false; h->mdc_CL = 0;
    const int K = h->md2_K, np = K * MD2S_T, GW = K * MD2S_NW;
    // ---- batch arrays ----
    CU(cudaMalloc((void**)&h->d_desc, (size_t)n_struct * sizeof(StructDesc)));
    CU(cudaMemcpyAsync(h->d_desc, h->desc.data(), (size_t)n_struct * sizeof(StructDesc), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMalloc((void**)&h->d_xyzr, na * sizeof(float4))); CU(cudaMalloc((void**)&h->d_xyzr0, na * sizeof(float4)));
    CU(cudaMalloc((void**)&h->d_kind, na)); CU(cudaMalloc((void**)&h->d_envmask, na * sizeof(uint32_t)));
    CU(cudaMalloc((void**)&h->d_bond_ptr, (size_t)n_struct * (n + 1) * sizeof(int32_t)));
    CU(cudaMalloc((void**)&h->d_bond_items, (size_t)std::max<int64_t>(2 * bonds, 1) * sizeof(uint32_t)));
    CU(cudaMalloc((void**)&h->d_btype_f, (size_t)btypes * sizeof(float2))); CU(cudaMalloc((void**)&h->d_btype_d, (size_t)btypes * sizeof(double2)));
    CU(cudaMalloc((void**)&h->d_asm_bl, (size_t)std::max<int64_t>(3 * bonds, 1) * sizeof(int32_t)));
    CU(cudaMalloc((void**)&h->d_asm_csr, (size_t)std::max<int64_t>(4 * bonds, 1) * sizeof(int32_t)));
    CU(cudaMalloc((void**)&h->d_asm_dec, (size_t)std::max<int64_t>(2 * bonds, 1)));
    CU(cudaMalloc((void**)&h->d_asm_atom, na * 3 * sizeof(uint16_t)));
    CU(cudaMalloc((void**)&h->d_md2_perm, (size_t)n_struct * np * sizeof(uint16_t)));
    CU(cudaMalloc((void**)&h->d_md2_eblk, (size_t)n_struct * GW * sizeof(int4)));
    CU(cudaMalloc((void**)&h->d_md2_desc, (size_t)n_struct * sizeof(Md2Desc)));
    CU(cudaMemsetAsync(h->d_asm_dec, 0, (size_t)std::max<int64_t>(2 * bonds, 1), h->stream));
    AsmOut O;
    memset(&O, 0, sizeof O);
    O.desc = h->d_desc; O.xyzr = h->d_xyzr; O.xyzr0 = h->d_xyzr0; O.kind = h->d_kind; O.envmask = h->d_envmask;
    O.bond_ptr = h->d_bond_ptr; O.bond_items = h->d_bond_items; O.btype_f = h->d_btype_f; O.btype_d = h->d_btype_d;
    O.bl_i = h->d_asm_bl; O.bl_j = h->d_asm_bl + bonds; O.bl_key = h->d_asm_bl + 2 * bonds;
    O.bidx = h->d_asm_csr; O.twin = h->d_asm_csr + 2 * bonds; O.dec = h->d_asm_dec;
    O.cntA = h->d_asm_atom; O.cntB = h->d_asm_atom + na; O.pos = h->d_asm_atom + 2 * na;
    O.perm = h->d_md2_perm; O.eblk = h->d_md2_eblk; O.ell_size = h->d_asm_cnt + 2 * n_struct; O.mdesc = h->d_md2_desc; O.K = K;
    const size_t sm_b = asm_bonds_smem(G, n), sm_o = asm_orient_smem(n), sm_l = asm_layout_smem(n, K);
    CU(cudaFuncSetAttribute(k_asm_bonds, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_b));
    CU(cudaFuncSetAttribute(k_asm_orient, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_o));
    CU(cudaFuncSetAttribute(k_asm_layout, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_l));
    k_asm_bonds<<<n_struct, ASM_T, sm_b, h->stream>>>(I, O);
    CU(cudaGetLastError());
    k_asm_orient<<<n_struct, 32, sm_o, h->stream>>>(I, O);
    CU(cudaGetLastError());
    k_asm_layout<<<n_struct, ASM_T, sm_l, h->stream>>>(I, O);
    CU(cudaGetLastError());
    std::vector<int32_t> ell((size_t)n_struct);
    CU(copy_on(h->stream, ell.data(), h->d_asm_cnt + 2 * n_struct, ell.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
    h->md2_desc.assign((size_t)n_struct, Md2Desc());
    int64_t ell_total = 0;
    for (int sl = 0; sl < n_struct; sl++) {
        Md2Desc& m = h->md2_desc[sl];
        m.ell_item_off = ell_total; m.perm_off = sl * np; m.ell_desc_off = sl * GW; m.K = K; m.pad = 0;
        ell_total += ell[sl];
    }
    CU(cudaMemcpyAsync(h->d_md2_desc, h->md2_desc.data(), (size_t)n_struct * sizeof(Md2Desc), cudaMemcpyHostToDevice, h->stream));
    CU(cudaMalloc((void**)&h->d_md2_ell_items, (size_t)std::max<int64_t>(ell_total, 1) * sizeof(uint32_t)));
    CU(cudaMalloc((void**)&h->d_md2_ell_act, (size_t)std::max<int64_t>(ell_total, 1) * sizeof(uint32_t)));
    O.ell_items = h->d_md2_ell_items;
    k_asm_ell<<<n_struct, ASM_T, 0, h->stream>>>(I, O);
    CU(cudaGetLastError());
    // ---- what upload() allocates besides ----
    h->xyzr.resize(na);
    CU(cudaMalloc((void**)&h->d_nl_cnt, na * sizeof(uint16_t)));
    CU(cudaMalloc((void**)&h->d_xd, na * 3 * sizeof(double))); CU(cudaMalloc((void**)&h->d_fd, na * 3 * sizeof(double)));
    CU(cudaMalloc((void**)&h->d_work, na * 12 * sizeof(double)));
    CU(cudaMalloc((void**)&h->d_res, (size_t)n_struct * sizeof(igmk_result_t))); CU(cudaMalloc((void**)&h->d_flags, (size_t)n_struct * sizeof(int32_t)));
    CU(cudaMemsetAsync(h->d_res, 0, (size_t)n_struct * sizeof(igmk_result_t), h->stream));
    CU(cudaMemsetAsync(h->d_flags, 0, (size_t)n_struct * sizeof(int32_t), h->stream));
    if (ctx_is_big<double>(n, false) || ctx_is_big<float>(n, false)) CU(cudaMalloc((void**)&h->d_big, (na + 32 * (size_t)n_struct) * 10 * sizeof(double)));
    CU(cudaMalloc((void**)&h->d_md2_gf, (size_t)n_struct * np * MD2_INBOX * sizeof(float4)));
    if (int rc = alloc_run_buffers(h)) return rc;
    h->dev_env_abck.assign(tmpl->env_abck, tmpl->env_abck + 4 * (size_t)tmpl->n_env);
    h->dev_env_off.assign(tmpl->env_off, tmpl->env_off + tmpl->n_env + 1);
    h->dev_env_mem.assign(tmpl->env_members, tmpl->env_members + tmpl->env_off[tmpl->n_env]);
    h->h2d_bytes = total + (size_t)n_struct * (sizeof(StructDesc) + sizeof(Md2Desc));
    h->dev_built = true; h->dirty = false; h->viol_valid = false;
    if (first_slot_out) *first_slot_out = 0;
    if (trace_sp)
        fprintf(stderr, "[igmk trace] submit_population_dev of %d: %.1f ms in all (%lld bonds, %lld sliced-ELL items)\n",
                n_struct, tnow() - tsp0, (long long)bonds, (long long)ell_total);
    return 0;
}

// One array of the staged batch as it sits on the device, for tests that compare the two assembly paths
// (0 desc, 1 xyzr, 2 xyzr0, 3 kind, 4 envmask, 5 bond_ptr, 6 bond_items, 7 btype_f, 8 btype_d, 9 md2_desc, 10 perm,
// 11 eblk, 12 ell_items).  *n_bytes = size of the array; the copy is made when it fits max_bytes.
extern "C" int igmk_debug_fetch_array(igmk_handle_t* h, int32_t which, void* out, int64_t max_bytes, int64_t* n_bytes)
{
    if (!h || !n_bytes) return fail("igmk_debug_fetch_array: NULL argument");
    if (h->device < 0) {                   // staging-only handle: the layout the upload would build, from the host vectors
        if (h->desc.empty()) return fail("igmk_debug_fetch_array: nothing submitted");
        if (h->md2_perm.empty() || h->dirty) {
            h->md2_K = md2_pick_K(h->max_atoms);
            if (h->md2_K == 0) return fail("igmk_debug_fetch_array: structure too large for the one-CTA layout");
            md2_build_owner(h);
            if (!h->md2_error.empty()) return fail("igmk: " + h->md2_error);
            h->dirty = false;
        }
        const void* src = nullptr; int64_t bytes = 0;
        switch (which) {
            case 0: src = h->desc.data(); bytes = h->desc.size() * sizeof(StructDesc); break;
            case 1: src = h->xyzr.data(); bytes = h->xyzr.size() * sizeof(float4); break;
            case 2: src = h->xyzr0.data(); bytes = h->xyzr0.size() * sizeof(float4); break;
            case 3: src = h->kind.data(); bytes = h->kind.size(); break;
            case 4: src = h->envmask.data(); bytes = h->envmask.size() * sizeof(uint32_t); break;
            case 5: src = h->bond_ptr.data(); bytes = h->bond_ptr.size() * sizeof(int32_t); break;
            case 6: src = h->bond_items.data(); bytes = h->bond_items.size() * sizeof(uint32_t); break;
            case 7: src = h->btype_f.data(); bytes = h->btype_f.size() * sizeof(float2); break;
            case 8: src = h->btype_d.data(); bytes = h->btype_d.size() * sizeof(double2); break;
            case 9: src = h->md2_desc.data(); bytes = h->md2_desc.size() * sizeof(Md2Desc); break;
            case 10: src = h->md2_perm.data(); bytes = h->md2_perm.size() * sizeof(uint16_t); break;
            case 11: src = h->md2_eblk.data(); bytes = h->md2_eblk.size() * sizeof(int4); break;
            case 12: src = h->md2_ell_items.data(); bytes = h->md2_ell_items.size() * sizeof(uint32_t); break;
            default: return fail("igmk_debug_fetch_array: no such array");
        }
        *n_bytes = bytes;
        if (out && bytes > 0 && bytes <= max_bytes) memcpy(out, src, (size_t)bytes);
        return 0;
    }
    CU(cudaSetDevice(h->device));
    if (int rc = upload(h)) return rc;
    const size_t ns = h->desc.size(), na = h->xyzr.size();
    if (ns == 0) return fail("igmk_debug_fetch_array: nothing submitted");
    const StructDesc& dl = h->desc.back();
    int64_t items = 0, btypes = 0, ell = 0;
    {
        int32_t last_ptr = 0;
        CU(copy_on(h->stream, &last_ptr, h->d_bond_ptr + dl.bond_ptr_off + dl.n_atoms, sizeof(int32_t), cudaMemcpyDeviceToHost));
        items = dl.bond_item_off + last_ptr;
        btypes = h->dev_built ? -1 : (int64_t)h->btype_f.size();
        if (btypes < 0) {            // last structure's offset + max(its types, 2): the count sits in d_asm_cnt
            int32_t nty = 0;
            CU(copy_on(h->stream, &nty, h->d_asm_cnt + ns + (ns - 1), sizeof(int32_t), cudaMemcpyDeviceToHost));
            btypes = dl.btype_off + std::max(nty, 2);
        }
        if (h->md2_K > 0 && !h->md_big && h->mdc_CL == 0) {
            const int GW = h->md2_K * MD2S_NW;
            int4 e;
            CU(copy_on(h->stream, &e, h->d_md2_eblk + ns * (size_t)GW - 1, sizeof(int4), cudaMemcpyDeviceToHost));
            ell = h->md2_desc.back().ell_item_off + e.z + (int64_t)e.w * 4 * 32;
        }
    }
    const void* src = nullptr; int64_t bytes = 0;
    const int K = h->md2_K;
    switch (which) {
        case 0: src = h->d_desc; bytes = ns * sizeof(StructDesc); break;
        case 1: src = h->d_xyzr; bytes = na * sizeof(float4); break;
        case 2: src = h->d_xyzr0; bytes = na * sizeof(float4); break;
        case 3: src = h->d_kind; bytes = na; break;
        case 4: src = h->d_envmask; bytes = na * sizeof(uint32_t); break;
        case 5: src = h->d_bond_ptr; bytes = (dl.bond_ptr_off + dl.n_atoms + 1) * sizeof(int32_t); break;
        case 6: src = h->d_bond_items; bytes = items * sizeof(uint32_t); break;
        case 7: src = h->d_btype_f; bytes = btypes * sizeof(float2); break;
        case 8: src = h->d_btype_d; bytes = btypes * sizeof(double2); break;
        case 9: src = h->d_md2_desc; bytes = ns * sizeof(Md2Desc); break;
        case 10: src = h->d_md2_perm; bytes = ns * (size_t)K * (h->mdc_CL > 0 ? MD2_T * h->mdc_CL : MD2S_T) * sizeof(uint16_t); break;
        case 11: src = h->d_md2_eblk; bytes = ns * (size_t)K * (h->mdc_CL > 0 ? MD2_NW * h->mdc_CL : MD2S_NW) * sizeof(int4); break;
        case 12: src = h->d_md2_ell_items; bytes = ell * sizeof(uint32_t); break;
        default: return fail("igmk_debug_fetch_array: no such array");
    }
    *n_bytes = bytes;
    if (out && bytes > 0 && bytes <= max_bytes) {
        if (!src) return fail("igmk_debug_fetch_array: array not allocated on this path");
        CU(copy_on(h->stream, out, src, (size_t)bytes, cudaMemcpyDeviceToHost));
    }
    return 0;
}

// =========================================================================================
// Single point for callers that must not allocate (see include/igmk.h).
//
// The regular igmk_single_point keeps its neighbour table, neighbour counts and the coordinates of
// the last neighbour build in cudaMalloc'ed buffers that grow with the batch.  This variant runs the
// same device code for ONE structure with that scratch in a fixed array that is part of the loaded
// module, so nothing is allocated, freed or resized at call time.
namespace {

constexpr int NOALLOC_NL_WIDTH = 96;
constexpr size_t NOALLOC_NL_BYTES = (size_t)IGMK_NOALLOC_MAX_ATOMS * NOALLOC_NL_WIDTH * sizeof(uint16_t);
constexpr size_t NOALLOC_CNT_BYTES = ((size_t)IGMK_NOALLOC_MAX_ATOMS * sizeof(uint16_t) + 255) / 256 * 256;
constexpr size_t NOALLOC_BUILD_BYTES = (size_t)IGMK_NOALLOC_MAX_ATOMS * 3 * sizeof(double);
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_NL_BYTES + NOALLOC_CNT_BYTES + NOALLOC_BUILD_BYTES;

// ~0.72 MB held by every GPU that loads the library, until the process exits.  In host code this
// name is only a handle for cudaGetSymbolAddress -- it is NOT a device address.
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};   // per device: device address of the array above
std::mutex g_noalloc_mutex;                                      // the first calls of several host threads set up once

// the entry point works on `device` and leaves the caller's current device as it found it
struct DeviceGuard {
    int prev = -1;
    cudaError_t enter(int device)
    {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) { prev = -1; return e; }
        return prev == device ? cudaSuccess : cudaSetDevice(device);
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

struct NoallocArgs {
    StructDesc desc;
    const float4* xyzr; const uint8_t* kind; const uint32_t* envmask;
    const int32_t* bond_ptr; const uint32_t* bond_items; const double2* btype;
    unsigned char* scratch;
    double* forces; double* energies; int32_t* status;
    double ev_scale, env_scale, bond_pref;
};

__global__ void __launch_bounds__(MD_THREADS, 1)
k_single_point_noalloc(const __grid_constant__ NoallocArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    typedef double T;
    const StructDesc* d = &a.desc;
    Batch B;
    B.n_struct = 1; B.max_atoms = d->n_atoms; B.desc = d;
    B.xyzr = nullptr; B.kind = a.kind; B.envmask = a.envmask;
    B.bond_ptr = a.bond_ptr; B.bond_items = a.bond_items; B.btype_f = nullptr; B.btype_d = a.btype;
    B.nl = (uint16_t*)a.scratch; B.nl_width = NOALLOC_NL_WIDTH;
    B.nl_cnt = (uint16_t*)(a.scratch + NOALLOC_NL_BYTES);
    B.xd = nullptr; B.fd = nullptr; B.work = nullptr; B.big = nullptr; B.res = nullptr; B.flags = nullptr;
    Ctx<T> c;
    ctx_carve<T>(c, smem, d->n_atoms, false, (T*)(a.scratch + NOALLOC_NL_BYTES + NOALLOC_CNT_BYTES));
    ctx_bind(c, B, d);
    c.set_stage(a.ev_scale, a.env_scale);
    c.bond_pref = a.bond_pref;
    for (int i = threadIdx.x; i < c.n; i += blockDim.x) {
        const float4 p = a.xyzr[i];
        double4 q; q.x = p.x; q.y = p.y; q.z = p.z; q.w = p.w;
        c.x[i] = q;
    }
    __syncthreads();
    c.rebuild();               // rewrites every neighbour row / count / build coordinate it later reads
    T e[10];
    c.template forces<true>(e);
    for (int i = threadIdx.x; i < c.n; i += blockDim.x) {
        a.forces[3 * i] = c.fx[i]; a.forces[3 * i + 1] = c.fy[i]; a.forces[3 * i + 2] = c.fz[i];
    }
    if (threadIdx.x == 0) {
        double tot = e[0] + e[1];
        for (int k = 0; k < IGMK_MAX_ENVELOPES; k++) {
            a.energies[3 + k] = k < d->n_env ? e[2 + k] : 0.0;
            if (k < d->n_env) tot += e[2 + k];
        }
        a.energies[0] = tot; a.energies[1] = e[0]; a.energies[2] = e[1];
        *a.status = c.overflow ? 1 : (isfinite(tot) ? 0 : 2);
    }
}

}  // namespace

extern "C" int igmk_single_point_noalloc(int device, void* stream, int32_t n_atoms,
                                         const float* d_xyzr, const uint8_t* d_kind, const uint32_t* d_envmask,
                                         const int32_t* d_bond_ptr, const uint32_t* d_bond_items, const double* d_btype,
                                         int32_t n_env, const double* env_abck, double evfactor, float max_radius,
                                         double ev_scale, double env_scale, double bond_prefactor,
                                         double* d_forces, double* d_energies, int32_t* d_status)
{
    if (!d_xyzr || !d_kind || !d_envmask || !d_bond_ptr || !d_bond_items || !d_btype || !d_forces || !d_energies || !d_status)
        return fail("igmk_single_point_noalloc: NULL argument");
    if (n_atoms < 1 || n_atoms > IGMK_NOALLOC_MAX_ATOMS) {
        char msg[200];
        snprintf(msg, sizeof msg, "igmk_single_point_noalloc: %d atoms; the built-in scratch array holds at most %d "
                 "(use igmk_single_point, which allocates what it needs)", n_atoms, IGMK_NOALLOC_MAX_ATOMS);
        return fail(msg);
    }
    if (n_env < 0 || n_env > IGMK_MAX_ENVELOPES || (n_env > 0 && !env_abck)) return fail("igmk_single_point_noalloc: bad envelope arguments");
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail("igmk_single_point_noalloc: bad device ordinal");
    DeviceGuard guard;
    CU(guard.enter(device));
    const size_t sm = ctx_smem_bytes<double>(n_atoms, false);
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[device]) {
            // one-time set-up on this device (with it current): resolve the array's device address
            void* p = nullptr;
            CU(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            CU(cudaFuncSetAttribute(k_single_point_noalloc, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    (int)ctx_smem_bytes<double>(IGMK_NOALLOC_MAX_ATOMS, false)));
            g_noalloc_ptr[device] = (unsigned char*)p;
        }
    }
    NoallocArgs a;
    memset(&a, 0, sizeof a);
    a.desc.n_atoms = n_atoms; a.desc.n_env = n_env; a.desc.evfactor = evfactor; a.desc.maxrad = max_radius;
    for (int e = 0; e < n_env; e++) for (int k = 0; k < 4; k++) a.desc.env[e][k] = env_abck[4 * e + k];
    a.xyzr = (const float4*)d_xyzr; a.kind = d_kind; a.envmask = d_envmask;
    a.bond_ptr = d_bond_ptr; a.bond_items = d_bond_items; a.btype = (const double2*)d_btype;
    a.scratch = g_noalloc_ptr[device];
    a.forces = d_forces; a.energies = d_energies; a.status = d_status;
    a.ev_scale = ev_scale; a.env_scale = env_scale; a.bond_pref = bond_prefactor;
    k_single_point_noalloc<<<1, MD_THREADS, sm, (cudaStream_t)stream>>>(a);
    CU(cudaGetLastError());
    return 0;
}
