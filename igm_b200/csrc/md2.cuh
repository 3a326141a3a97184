// md2.cuh -- the annealing-MD kernel of the M-step (lammps.py:285-351), one CTA per structure.
//
// Work decomposition: THREAD PER ATOM.  Thread t owns the atoms at positions t, t+1024, ...
// (K slots per thread) of a host-chosen ordering in which the 32 atoms of a warp-slot have
// similar bond counts.  Each thread gathers the complete force on its own atoms (no atomics, no
// cross-lane reductions => bit-reproducible), and keeps their velocities and forces in registers.
// Shared memory holds only what other threads read: the positions (float4 x,y,z,radius), a copy of
// them in cell order for the neighbour search, the coordinates of the last neighbour build and the
// cell grid.  Lists are laid out so that a warp reads them with full 32-lane coalescing:
//   * neighbour table   nl[batch][position][4]   (uint16; 4 neighbours per 8-byte load; rebuilt in-kernel)
//   * bond items        ell[warp-slot][batch][lane][4] (uint32 partner | type<<16; 4 items per 16-byte load)
//   * bond candidates   same layout, rewritten by the first force evaluation after every rebuild
// They live in global memory and are streamed through L1/L2.  Rows are padded with self-references
// (r = 0, no force) so that the inner loops carry no bounds tests.
#pragma once
#include "mstep.cuh"

namespace igmk {

#ifndef MD2_THREADS
#define MD2_THREADS 1024
#endif
constexpr int MD2_T = MD2_THREADS;    // threads per CTA
constexpr int MD2_NW = MD2_T / 32;    // warps per CTA
constexpr int MD2_KMAX = 8;           // slots per thread supported (n_atoms <= 8192 on the 1-CTA path)
constexpr int MD2_NLW0 = 256;         // initial neighbour-table width (columns)
constexpr uint16_t MD2_NONE = 0xffffu;

struct Md2View {
    const uint16_t* perm;      // [sum np] position -> original atom (0xffff: empty position)
    const int32_t*  ell_off;   // per structure: 32*K+1 offsets (in items) of each warp-slot block
    const uint32_t* ell_items; // sliced-ELL bond items
    uint32_t*       ell_act;   // same size: the items that can be active before the next rebuild (k_md2 scratch)
    uint16_t*       nl;        // [sum np * nl_width] column-major per structure
    int32_t         nl_width;
    long long*      prof;      // optional [n_struct][8] SM cycles (IGMK_PROF=1): force loops, rebuild search, rebuilds,
                               // number of rebuilds, then cycles of annealing stages 0..3
};

struct Md2Desc {               // per structure
    int64_t ell_item_off;
    int32_t perm_off;          // offset into perm (== sum of np of the structures before)
    int32_t ell_desc_off;      // offset into ell_off
    int32_t K;
    int32_t pad;
};

// single-instruction approximations (MUFU) without the denormal fix-up code of the CUDA intrinsics;
// their ~1e-6 relative error is below the float32 round-off the MD trajectory carries anyway
__device__ __forceinline__ float fast_rsqrt(float v) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float fast_rcp(float v) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float fast_sin(float v) { float r; asm("sin.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
// shared-memory load through an explicit 32-bit shared address (keeps the address arithmetic to one shift-add)
__device__ __forceinline__ float4 lds_f4(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}

__host__ __device__ inline size_t md2_smem_bytes(int K)
{
    const size_t np = (size_t)K * MD2_T;
    size_t b = np * 16;                       // x
    b += np * 16;                             // positions in cell order (neighbour search)
    b += np * 12;                             // build coordinates
    b += ((size_t)MAX_CELLS + 1) * 4;         // cell_start
    b += np * 2 * 2;                          // cell_atoms, cell_of
    b += 2 * 32 * 8 * 4;                      // reduction scratch
    b += IGMK_MAX_ENVELOPES * 4 * 4;          // envelope parameters
    b += (size_t)K * MD2_NW * 4;              // active bond batches per warp-slot
    return (b + 255) / 256 * 256;
}

template <int K>
__global__ void __launch_bounds__(MD2_T, 1)
k_md2(Batch B, Md2View V, const Md2Desc* __restrict__ mdesc, MdParams P)
{
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int np = K * MD2_T;
    const StructDesc* d = B.desc + blockIdx.x;
    const Md2Desc md = mdesc[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = d->n_atoms;

    float4* x = (float4*)smem;
    float4* xsort = x + np;
    float* xbx = (float*)(xsort + np); float* xby = xbx + np; float* xbz = xby + np;
    uint32_t* cell_start = (uint32_t*)(xbz + np);
    uint16_t* cell_atoms = (uint16_t*)(cell_start + MAX_CELLS + 1);
    uint16_t* cell_of = cell_atoms + np;
    float* rbuf = (float*)(cell_of + np);
    float* envp = rbuf + 2 * 256;
    int* bact = (int*)(envp + IGMK_MAX_ENVELOPES * 4);      // [K][32] batches of candidate bond items per warp-slot
    int rb_par = 0;

    const igmk_protocol_t& pr = P.prot;
    const uint16_t* perm = V.perm + md.perm_off;
    uint32_t* eact = V.ell_act + md.ell_item_off;
    const int32_t* eoff = V.ell_off + md.ell_desc_off;
    const uint32_t* eitems = V.ell_items + md.ell_item_off;
    uint16_t* nl = V.nl + (size_t)md.perm_off * V.nl_width;
    const int nlw = V.nl_width;
    const float2* btype = B.btype_f + d->btype_off;
    float4* xg = B.xyzr + d->atom_off;

    // ---- per-slot state in registers ----
    int   kind[K];            // IGMK_KIND_*, 3 = empty position
    uint32_t emask[K];
    int   ncnt[K];            // neighbours in my row
    int   boff[K], bcols[K];  // my warp-slot's block of bond items
    float vx[K], vy[K], vz[K], fx[K], fy[K], fz[K];
#pragma unroll
    for (int k = 0; k < K; k++) {
        const int p = k * MD2_T + tid;
        const uint16_t a = perm[p];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        kind[k] = 3; emask[k] = 0;
        if (a != MD2_NONE) {
            q = xg[a];
            kind[k] = B.kind[d->atom_off + a];
            emask[k] = B.envmask[d->atom_off + a];
        }
        x[p] = q;
        const int g = k * MD2_NW + warp;
        boff[k] = eoff[g];
        bcols[k] = (eoff[g + 1] - eoff[g]) >> 5;
        vx[k] = vy[k] = vz[k] = 0.f; fx[k] = fy[k] = fz[k] = 0.f; ncnt[k] = 0;
    }
    // the two most frequent bond types stay in registers (the host orders the type table by use)
    const float2 bt0 = btype[0], bt1 = btype[1];
    const float skin = d->maxrad * P.skin_factor;
    const float lim = 0.25f * skin * skin;
    const float bond_pref2 = 2.f * (float)pr.bond_prefactor;
    float ev_c = -1.f;                 // evfactor * stage scale / pi  (force prefactor / rc), < 0: no sterics
    int n_rebuild = 0, overflow = 0;
    bool bond_build = true;
    long long t_search = 0;
    __syncthreads();

    // ------------------------------------------------------------------------------------
    auto block_sum1 = [&](float v) -> float {
        float* buf = rbuf + (rb_par ^= 1) * 256;
        v = warp_sum(v);
        if (lane == 0) buf[warp] = v;
        __syncthreads();
        return warp_sum(lane < MD2_NW ? buf[lane] : 0.f);     // <= 32 partials: one per lane, same tree in every warp
    };

    auto set_stage = [&](double ev_scale, double env_scale) {
        ev_c = d->evfactor < 0 ? -1.f : (float)(d->evfactor * ev_scale / 3.14159265358979323846);
        if (tid < IGMK_MAX_ENVELOPES * 4) {
            const int e = tid >> 2, c = tid & 3;
            envp[tid] = e < d->n_env ? (float)(c < 3 ? d->env[e][c] * env_scale : d->env[e][3]) : 0.f;
        }
        __syncthreads();
    };

    const uint32_t xs = (uint32_t)__cvta_generic_to_shared(x);
    // ---- neighbour table: cell binning in shared memory, every thread fills its own rows ----
    auto rebuild = [&]() {
        float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const float4 q = x[p];
            xbx[p] = q.x; xby[p] = q.y; xbz[p] = q.z;
            if (kind[k] == IGMK_KIND_BEAD) {
                lo[0] = fminf(lo[0], q.x); hi[0] = fmaxf(hi[0], q.x);
                lo[1] = fminf(lo[1], q.y); hi[1] = fmaxf(hi[1], q.y);
                lo[2] = fminf(lo[2], q.z); hi[2] = fmaxf(hi[2], q.z);
            }
        }
        float* rb = rbuf + (rb_par ^= 1) * 256;
#pragma unroll
        for (int c = 0; c < 3; c++) { lo[c] = warp_min(lo[c]); hi[c] = warp_max(hi[c]); }
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < 3; c++) { rb[warp * 6 + c] = lo[c]; rb[warp * 6 + 3 + c] = hi[c]; }
        }
        for (int c = tid; c <= MAX_CELLS; c += MD2_T) cell_start[c] = 0;
        __syncthreads();
#pragma unroll
        for (int c = 0; c < 3; c++) {
            lo[c] = warp_min(lane < MD2_NW ? rb[lane * 6 + c] : 1e30f); hi[c] = warp_max(lane < MD2_NW ? rb[lane * 6 + 3 + c] : -1e30f);
        }
        const float rlist = 2.f * d->maxrad + skin;
        int dim[3]; float inv[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float ext = hi[c] - lo[c];
            if (!(ext > 0.f)) ext = 0.f;
            int m = (int)(ext / rlist) + 1;
            if (m > MAX_CELL_DIM) m = MAX_CELL_DIM;
            float cw = ext / (float)m;
            if (cw < rlist) cw = rlist;
            dim[c] = m; inv[c] = 1.f / cw;
        }
        const int ncell = dim[0] * dim[1] * dim[2];
        // counts per cell (integer atomics: the counts do not depend on the order)
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            int c = MD2_NONE;
            if (kind[k] == IGMK_KIND_BEAD) {
                const float4 q = x[p];
                int cx = (int)((q.x - lo[0]) * inv[0]), cy = (int)((q.y - lo[1]) * inv[1]), cz = (int)((q.z - lo[2]) * inv[2]);
                cx = min(max(cx, 0), dim[0] - 1); cy = min(max(cy, 0), dim[1] - 1); cz = min(max(cz, 0), dim[2] - 1);
                c = (cz * dim[1] + cy) * dim[0] + cx;
                atomicAdd(&cell_start[c + 1], 1u);
            }
            cell_of[p] = (uint16_t)c;
        }
        __syncthreads();
        // inclusive scan of cell_start[1..ncell] (cell_start[c+1] held the count of cell c)
        {
            const int per = (ncell + MD2_T - 1) / MD2_T;
            const int c0 = min(tid * per, ncell), c1 = min(c0 + per, ncell);
            int s = 0;
            for (int c = c0; c < c1; c++) s += cell_start[c + 1];
            int incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            int* ib = (int*)(rbuf + (rb_par ^= 1) * 256);
            if (lane == 31) ib[warp] = incl;
            __syncthreads();
            int base = 0;
            for (int w = 0; w < warp; w++) base += ib[w];
            int run = base + incl - s;
            for (int c = c0; c < c1; c++) { const int cnt = cell_start[c + 1]; run += cnt; cell_start[c + 1] = run; }
        }
        __syncthreads();
        // scatter with a per-cell cursor kept in cell_of's upper half?  no: reuse the counts trick --
        // drop atoms unordered behind the cell start, then sort each (tiny) cell by index so that the
        // row order, and with it the floating-point summation order, is independent of the race.
        uint32_t* cursor = (uint32_t*)xbx;          // aliases the build x, which is re-written below
        for (int c = tid; c < ncell; c += MD2_T) cursor[c] = 0;
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const int c = cell_of[p];
            if (c != MD2_NONE) {
                const uint32_t pos = atomicAdd(&cursor[c], 1u);
                cell_atoms[cell_start[c] + pos] = (uint16_t)p;
            }
        }
        __syncthreads();
        for (int c = tid; c < ncell; c += MD2_T) {
            const int b = cell_start[c], e = cell_start[c + 1];
            for (int q = b + 1; q < e; q++) {
                const uint16_t v = cell_atoms[q]; int r = q - 1;
                while (r >= b && cell_atoms[r] > v) { cell_atoms[r + 1] = cell_atoms[r]; r--; }
                cell_atoms[r + 1] = v;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) xbx[k * MD2_T + tid] = x[k * MD2_T + tid].x;
        // positions in cell order: the search below walks them contiguously
        const long long ts0 = clock64();
        const int nbead = cell_start[ncell];
        for (int s0 = tid; s0 < nbead; s0 += MD2_T) xsort[s0] = x[cell_atoms[s0]];
        __syncthreads();
        // neighbour search.  Thread t takes the t-th bead IN CELL ORDER (not one of its own atoms), so the
        // lanes of a warp sit in the same or adjacent cells, scan nearly the same ranges (little divergence,
        // broadcast shared-memory reads) and need no index indirection until a candidate is accepted.
        // Row layout: nl[batch][position][4] -- the force loop fetches 4 neighbours with one 8-byte load.
        int ovf = 0;
        for (int s0 = tid; s0 < nbead; s0 += MD2_T) {
            const int p = cell_atoms[s0];
            const float4 q = xsort[s0];
            const int c = cell_of[p];
            const int cx = c % dim[0], cy = (c / dim[0]) % dim[1], cz = c / (dim[0] * dim[1]);
            const int ax0 = max(cx - 1, 0), ax1 = min(cx + 1, dim[0] - 1);
            uint16_t* row = nl + ((size_t)p << 2);
            int cnt = 0;
            for (int dz = -1; dz <= 1; dz++) {
                const int az = cz + dz;
                if (az < 0 || az >= dim[2]) continue;
                for (int dy = -1; dy <= 1; dy++) {
                    const int ay = cy + dy;
                    if (ay < 0 || ay >= dim[1]) continue;
                    const int cb = (az * dim[1] + ay) * dim[0];
                    const int b = cell_start[cb + ax0], e = cell_start[cb + ax1 + 1];   // 3 cells, contiguous
                    for (int s1 = b; s1 < e; s1 += 4) {
                        bool hit[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            const int s2 = min(s1 + u, e - 1);
                            const float4 pj = xsort[s2];
                            const float ddx = q.x - pj.x, ddy = q.y - pj.y, ddz = q.z - pj.z;
                            const float rl = q.w + pj.w + skin;
                            hit[u] = (s1 + u < e) && (s2 != s0) && (ddx * ddx + ddy * ddy + ddz * ddz < rl * rl);
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            if (hit[u]) {
                                if (cnt < nlw) row[(((size_t)(cnt >> 2) * np) << 2) + (cnt & 3)] = cell_atoms[s1 + u];
                                cnt++;
                            }
                        }
                    }
                }
            }
            if (cnt > nlw) { ovf = 1; cnt = nlw; }
            for (int cc = cnt; cc & 3; cc++) row[(((size_t)(cc >> 2) * np) << 2) + (cc & 3)] = (uint16_t)p;   // self padding
            cell_of[p] = (uint16_t)cnt;            // hand the row length to the atom's owner
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) {
            const uint16_t c = cell_of[k * MD2_T + tid];
            ncnt[k] = c == MD2_NONE ? 0 : (int)c;
        }
        overflow |= __syncthreads_or(ovf);
        t_search += clock64() - ts0;
        bond_build = true;        // the next force evaluation also refreshes the bond candidate list
        n_rebuild++;
    };

    // ---- forces on my atoms (gather) -------------------------------------------------------
    // Rows of the neighbour table are padded with the atom's own index to a multiple of 4 and the bond
    // blocks with self-bonds of type 0: a self pair has r2 = 0 and contributes nothing, so the inner
    // loops need no bounds or validity tests and run branch-free.
    auto forces = [&]() {
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const float4 pi = x[p];
            float ax = 0.f, ay = 0.f, az = 0.f;
            // pair_style soft: F = A pi/rc sin(pi r/rc) r_hat, A = rc^2 ev/pi^2  =>  F/r = ev_c rc sin(pi r/rc) / r
            if (ev_c >= 0.f) {
                const uint2* col = (const uint2*)nl + p;
                const int nb4 = (ncnt[k] + 3) >> 2;
                for (int q = 0; q < nb4; q++) {
                    const uint2 w = col[(size_t)q * np];
                    const uint32_t j[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float4 pj = lds_f4(xs + (j[u] << 4));
                        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                        const float rc = pi.w + pj.w;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float rinv = fast_rsqrt(r2);
                        const float s = fast_sin(3.14159265358979323846f * (r2 * rinv) * fast_rcp(rc));
                        float fp = ev_c * rc * s * rinv;
                        fp = (r2 < rc * rc && r2 > 0.f) ? fp : 0.f;
                        ax += dx * fp; ay += dy * fp; az += dz * fp;
                    }
                }
            }
            // harmonic upper / lower bounds: F = -2 pref k (r - r0) r_hat when violated
            {
                // Bond candidates.  No atom moves more than skin/2 before the next rebuild, so no pair distance
                // changes by more than skin: an upper bound with r < r0 - skin (a lower bound with r > r0 + skin)
                // stays inactive until then.  The first evaluation after a rebuild walks the full list and
                // writes the candidates (original order kept) to `eact`; the others walk only those.  In a
                // converged population most restraints are satisfied with room to spare.
                const bool build = bond_build;
                const uint4* it = (const uint4*)((build ? eitems : eact) + boff[k]) + lane;   // [batch][lane][4 items]
                const int nb4 = build ? (bcols[k] >> 2) : bact[k * MD2_NW + warp];
                uint32_t* out = eact + boff[k] + (lane << 2);
                int ccnt = 0;
                uint4 nxt = nb4 > 0 ? it[0] : make_uint4(0, 0, 0, 0);
                for (int c = 0; c < nb4; c++) {
                    const uint4 cur = nxt;
                    if (c + 1 < nb4) nxt = it[(c + 1) * 32];                  // next batch in flight while this one is used
                    const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const uint32_t ti = v[u] >> 16;
                        float2 t = ti == 0 ? bt0 : bt1;                    // (r0, k); k < 0: lower bound
                        if (ti > 1) t = btype[ti];
                        const float4 pj = lds_f4(xs + ((v[u] & 0xffffu) << 4));
                        const float dx = pi.x - pj.x, dy = pi.y - pj.y, dz = pi.z - pj.z;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float r02 = t.x * t.x;
                        const bool act = (t.y > 0.f ? (r2 > r02) : (r2 < r02)) && r2 > 0.f;
                        float fb = -bond_pref2 * fabsf(t.y) * (1.f - t.x * fast_rsqrt(r2));
                        fb = act ? fb : 0.f;
                        ax += dx * fb; ay += dy * fb; az += dz * fb;
                        if (build) {
                            const float lo = fmaxf(t.x - skin, 0.f), hi = t.x + skin;
                            if (t.y > 0.f ? (r2 > lo * lo) : (r2 < hi * hi)) { out[((ccnt >> 2) << 7) + (ccnt & 3)] = v[u]; ccnt++; }
                        }
                    }
                }
                if (build) {
                    const int most = warp_max((ccnt + 3) >> 2);
                    for (int cc = ccnt; cc < (most << 2); cc++) out[((cc >> 2) << 7) + (cc & 3)] = (uint32_t)p;   // self-bond padding
                    if (lane == 0) bact[k * MD2_NW + warp] = most;
                }
            }
            // fix ellipsoidalenvelope
            uint32_t m = emask[k];
            while (m) {
                const int e = __ffs(m) - 1; m &= m - 1;
                const float ia = fast_rcp(envp[4 * e] - pi.w), ib = fast_rcp(envp[4 * e + 1] - pi.w),
                            ic = fast_rcp(envp[4 * e + 2] - pi.w), kk = envp[4 * e + 3];
                const float ux = pi.x * ia, uy = pi.y * ib, uz = pi.z * ic;
                const float k2 = ux * ux + uy * uy + uz * uz;
                const bool act = (kk > 0.f ? (k2 > 1.f) : (kk < 0.f && k2 < 1.f)) && k2 > 0.f;
                const float g = act ? fabsf(kk) * (1.f - fast_rsqrt(k2)) : 0.f;
                ax -= g * pi.x; ay -= g * pi.y; az -= g * pi.z;
            }
            if (kind[k] != IGMK_KIND_BEAD && kind[k] != IGMK_KIND_CENTROID) { ax = 0.f; ay = 0.f; az = 0.f; }   // frozen / empty
            fx[k] = ax; fy[k] = ay; fz[k] = az;
        }
        bond_build = false;
    };

    // ------------------------------------------------------------------------------------
    long long t_force = 0, t_integ = 0, t_build = 0;
    const long long t_begin = clock64();
    rebuild();
    t_build += clock64() - t_begin;
    const float dt = (float)pr.timestep, dtf = 0.5f * dt;
    const float dof = 3.f * (float)d->n_nonfixed - 3.f;
    float last_temp = 0.f;

    long long t_stage[4] = {0, 0, 0, 0};
    for (int st = 0; st < pr.n_stage; st++) {
        const long long ts_stage = clock64();
        set_stage(pr.evfactor[st], pr.envfactor[st]);
        for (int sub = (pr.relax_mdsteps > 0 ? 0 : 1); sub < 2; sub++) {
            const int nsteps = sub == 0 ? pr.relax_mdsteps : pr.mdsteps[st];
            const float t0 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstart[st];
            const float t1 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstop[st];
            const float xmax = sub == 0 ? (float)pr.relax_max_velocity : (float)pr.max_velocity;
            const float vlim2 = (xmax / dt) * (xmax / dt);
            // ---- velocity nonfixed create t0 seed+st (dist uniform, mom yes, loop all) ----
            {
                uint32_t seed = (uint32_t)((d->seed + st) % 2147483647ll);
                if (seed == 0) seed = 1;
                float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    float a = 0.f, b = 0.f, c = 0.f;
                    if (kind[k] == IGMK_KIND_BEAD || kind[k] == IGMK_KIND_CENTROID) {
                        uint32_t r = ranpark_jump(seed, 3u * (uint32_t)perm[k * MD2_T + tid] + 1u);
                        const double AM = 1.0 / 2147483647.0;
                        a = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); b = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); c = (float)(AM * r - 0.5);
                    }
                    vx[k] = a; vy[k] = b; vz[k] = c;
                    s0 += a; s1 += b; s2 += c;
                }
                s0 = block_sum1(s0); s1 = block_sum1(s1); s2 = block_sum1(s2);
                const float inv = d->n_nonfixed > 0 ? 1.f / (float)d->n_nonfixed : 0.f;
                float ke = 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    if (kind[k] == IGMK_KIND_BEAD || kind[k] == IGMK_KIND_CENTROID) {
                        vx[k] -= s0 * inv; vy[k] -= s1 * inv; vz[k] -= s2 * inv;
                        ke += vx[k] * vx[k] + vy[k] * vy[k] + vz[k] * vz[k];
                    }
                }
                ke = block_sum1(ke);
                const float tcur = dof > 0.f ? ke / dof : 0.f;
                const float fac = tcur > 0.f ? sqrtf(t0 / tcur) : 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) { vx[k] *= fac; vy[k] *= fac; vz[k] *= fac; }
            }
            // ---- run nsteps: velocity Verlet with the nve/limit speed cap + temp/rescale ----
            forces();                                   // run setup
            float scale = 1.f;
            for (int step = 0; step <= nsteps; step++) {
                const long long tc0 = clock64();
                if (step > 0) {
                    forces();
                    t_force += clock64() - tc0;
                    float ke = 0.f;
#pragma unroll
                    for (int k = 0; k < K; k++) {       // second half kick (frozen/empty: f = v = 0)
                        float a = vx[k] + dtf * fx[k], b = vy[k] + dtf * fy[k], c = vz[k] + dtf * fz[k];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[k] = a; vy[k] = b; vz[k] = c;
                        ke += a * a + b * b + c * c;
                    }
                    ke = block_sum1(ke);                // its barrier also ends every read of x for this step
                    const float tcur = dof > 0.f ? ke / dof : 0.f;
                    const float ttar = t0 + ((float)step / (float)nsteps) * (t1 - t0);
                    scale = 1.f;
                    if (tcur > 0.f && fabsf(tcur - ttar) > 0.1f) scale = sqrtf(ttar / tcur);
                    last_temp = tcur * scale * scale;
                } else {
                    __syncthreads();
                }
                if (step == nsteps) {
#pragma unroll
                    for (int k = 0; k < K; k++) { vx[k] *= scale; vy[k] *= scale; vz[k] *= scale; }
                    break;
                }
                int need = 0;
#pragma unroll
                for (int k = 0; k < K; k++) {           // rescale + first half kick + drift
                    if (kind[k] == IGMK_KIND_BEAD || kind[k] == IGMK_KIND_CENTROID) {
                        const int p = k * MD2_T + tid;
                        float a = vx[k] * scale + dtf * fx[k], b = vy[k] * scale + dtf * fy[k], c = vz[k] * scale + dtf * fz[k];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[k] = a; vy[k] = b; vz[k] = c;
                        float4 q = x[p];
                        q.x += dt * a; q.y += dt * b; q.z += dt * c;
                        x[p] = q;
                        const float ddx = q.x - xbx[p], ddy = q.y - xby[p], ddz = q.z - xbz[p];
                        need |= (ddx * ddx + ddy * ddy + ddz * ddz > lim);
                    }
                }
                const bool stale = __syncthreads_or(need);
                const long long tc1 = clock64();
                if (stale) { rebuild(); t_build += clock64() - tc1; }
                t_integ += tc1 - tc0;
            }
        }
        if (st < 4) t_stage[st] = clock64() - ts_stage;
    }
    // results: coordinates back in the original atom order (float32 like Particle.pos)
    int bad = 0;
#pragma unroll
    for (int k = 0; k < K; k++) {
        const uint16_t a = perm[k * MD2_T + tid];
        if (a != MD2_NONE) {
            const float4 q = x[k * MD2_T + tid];
            xg[a] = q;
            bad |= !(isfinite(q.x) && isfinite(q.y) && isfinite(q.z));
        }
    }
    bad = __syncthreads_or(bad);
    if (tid == 0) {
        igmk_result_t* r = B.res + blockIdx.x;
        const double dof_all = 3.0 * n - 3.0;     // thermo `temp` is over group all
        r->temperature = dof_all > 0 ? (double)last_temp * (double)dof / dof_all : 0.0;
        r->n_rebuilds = n_rebuild;
        B.flags[blockIdx.x] = (overflow ? 1 : 0) | (bad ? 2 : 0);
        if (V.prof) {
            long long* o = V.prof + 8 * (size_t)blockIdx.x;
            o[0] = t_force; o[1] = t_search; o[2] = t_build; o[3] = n_rebuild;
            for (int k = 0; k < 4; k++) o[4 + k] = t_stage[k];
        }
    }
}

}  // namespace igmk
