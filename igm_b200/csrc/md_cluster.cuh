// md_cluster.cuh -- annealing MD of ONE structure on a thread-block CLUSTER (north_star: "a cluster with distributed
// shared memory at 200 kb"; BASELINE configs[3]: 29 838 beads).
//
// The CL CTAs of a cluster split the atoms of the structure: CTA r owns positions r*NPC .. (r+1)*NPC-1 of a host-chosen
// ordering that follows the genome (a chromosome territory is compact in space, so most partners of an atom live in
// the same CTA) and, inside a CTA, sorts by owned bonds like k_md2.  Every CTA keeps ITS atoms' positions in its own
// shared memory and their velocities / forces in registers; a partner position is read through distributed shared
// memory (mapa + ld.shared::cluster: the owning CTA is encoded in the index).  Oriented bonds (owner / receiver, inbox
// flag + mailbox as in k_md2) exist only between atoms of ONE CTA; the host lists a bond between two CTAs at both ends
// (SYM), so inboxes and mailboxes are CTA-local and the barrier before the receive is a __syncthreads.  (The
// st.shared::cluster path for a remote inbox flag is still there; the layout no longer produces such items.)  The loops,
// the arithmetic and the list layouts are those of md2.cuh (thread per atom, oriented bonds, candidate lists); what
// changes is the scope of the synchronisation -- two cluster barriers per step (kinetic energy summed = every CTA has
// read the positions; positions written + rebuild flag) and one CTA barrier instead of three CTA barriers -- and the
// neighbour rebuild, which needs
// the whole structure: the cluster bins all beads into ONE cell grid in global memory (counts, scan by CTA 0,
// scatter, per-cell sort, cell-ordered copy of the positions), then every CTA searches its share of the beads IN CELL
// ORDER (not its own atoms: even load, lanes of a warp scan the same cells so the reads hit L1) and writes their
// rows.  Everything a cluster touches stays on its own GPU.
#pragma once
#include "md2.cuh"

namespace igmk {

constexpr int MDC_MAXC_DIM = 32;
constexpr int MDC_MAXC = MDC_MAXC_DIM * MDC_MAXC_DIM * MDC_MAXC_DIM;     // cells of the global grid
constexpr int MDC_MAX_CL = 16;
// bond item: bits 0-3 inbox slot q, 4-19 partner position (rank * NPC + offset), 20 SYM, 21-31 type
constexpr uint32_t MDC_IT_SYM = 1u << 20;
constexpr int MDC_IT_TSHIFT = 21;
constexpr int MDC_MAX_TYPES = 1 << 11;

struct MdcView {
    const uint16_t* perm;      // [n_struct][CL*NPC] position -> original atom (0xffff: empty)
    const uint32_t* ell_items; // per structure: blocks of warp-slot g = (rank*K + k)*32 + warp, layout as in md2.cuh
    uint32_t*       ell_act;
    const int4*     eblk;      // [n_struct][CL*K*32] {offset A, batches A, offset B, batches B}
    uint16_t*       nl;        // [n_struct][CL*NPC * nl_width] rows by global position, [batch][position][4]
    int32_t         nl_width;
    float4*         gf;        // mailboxes [n_struct][CL*NPC*16]
    // rebuild workspace in global memory, per structure
    float4*         xall;      // [CL*NPC] positions by global position
    float4*         xsorted;   // [CL*NPC] positions in cell order
    uint32_t*       sorted;    // [CL*NPC] global position of the s-th bead in cell order
    uint32_t*       cells;     // [2*(MAXC+1)] cell_start | cursor
    uint16_t*       rowlen;    // [CL*NPC]
    float*          bbox;      // [MDC_MAX_CL*8]
    long long*      prof;
    double*         dbg_forces;
};

__host__ __device__ inline size_t mdc_smem_bytes(int K)
{
    const size_t np = (size_t)K * MD2_T;
    size_t b = np * 16;                       // x
    b += np * 16;                             // inboxes
    b += np * 12;                             // build coordinates
    b += 2 * 32 * 8 * 4;                      // reduction scratch
    b += 2 * MDC_MAX_CL * 4 * 4;              // cluster exchange slots (two parities x up to 4 values per CTA)
    b += IGMK_MAX_ENVELOPES * 4 * 4;          // envelope parameters
    b += (size_t)K * MD2_NW * 16;             // bond block descriptors
    b += (size_t)K * MD2_NW * 8;              // candidate batches per warp-slot (A, B)
    b += 16 * 16;                             // compaction table of the neighbour search
    return (b + 255) / 256 * 256;
}

__device__ __forceinline__ void mdc_sync()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mdc_map(uint32_t local_shared_addr, uint32_t rank)
{
    uint32_t ra;
    asm("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_shared_addr), "r"(rank));
    return ra;
}
__device__ __forceinline__ float4 mdc_ld_f4(uint32_t cluster_addr)
{
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(cluster_addr));
    return v;
}
__device__ __forceinline__ void mdc_st_u8(uint32_t cluster_addr, uint32_t v)
{
    asm volatile("st.shared::cluster.u8 [%0], %1;" :: "r"(cluster_addr), "r"(v) : "memory");
}
__device__ __forceinline__ void mdc_st_f32(uint32_t cluster_addr, float v)
{
    asm volatile("st.shared::cluster.f32 [%0], %1;" :: "r"(cluster_addr), "f"(v) : "memory");
}

template <int K, int CL, bool PROF = false>
__global__ void __launch_bounds__(MD2_T, 1)
k_md2c(Batch B, MdcView V, const Md2Desc* __restrict__ mdesc, MdParams P)
{
    extern __shared__ __align__(16) unsigned char smem[];
    constexpr int np = K * MD2_T;                 // positions per CTA (a power of two)
    constexpr int NPS = np * CL;                  // positions per structure
    constexpr int LOGNP = K == 1 ? 10 : K == 2 ? 11 : 12;
    static_assert(K == 1 || K == 2 || K == 4, "positions per CTA must be a power of two");
    const int sidx = blockIdx.x / CL;
    uint32_t rank;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const StructDesc* d = B.desc + sidx;
    const Md2Desc md = mdesc[sidx];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = d->n_atoms;

    float4* x = (float4*)smem;
    uint4* inbox = (uint4*)(x + np);
    float* xbx = (float*)(inbox + np); float* xby = xbx + np; float* xbz = xby + np;
    float* rbuf = xbz + np;
    float* cbuf = rbuf + 2 * 256;                 // [2][MDC_MAX_CL][4] values other CTAs write for me
    float* envp = cbuf + 2 * MDC_MAX_CL * 4;
    int4* eblk = (int4*)(envp + IGMK_MAX_ENVELOPES * 4);
    int2* bact = (int2*)(eblk + K * MD2_NW);
    uint4* lut = (uint4*)(bact + K * MD2_NW);     // hit mask of four candidates -> byte selectors that pack the hits (as in k_md2)
    if (threadIdx.x < 16) {
        uint32_t sel[4] = {0u, 0u, 0u, 0u};
        int nsel = 0;
        for (int j = 0; j < 4; j++) if (threadIdx.x & (1 << j)) sel[nsel++] = (uint32_t)(2 * j) | ((uint32_t)(2 * j + 1) << 4);
        lut[threadIdx.x] = make_uint4(sel[0] | (sel[1] << 8), sel[2] | (sel[3] << 8),
                                      nsel >= 2 ? 0xffffffffu : (nsel == 1 ? 0xffffu : 0u), nsel >= 4 ? 0xffffffffu : (nsel == 3 ? 0xffffu : 0u));
    }
    int rb_par = 0, cb_par = 0;

    const igmk_protocol_t& pr = P.prot;
    const size_t sbase = (size_t)sidx * NPS;      // this structure's block in the per-position arrays
    const uint16_t* perm = V.perm + sbase + (size_t)rank * np;
    uint32_t* eact = V.ell_act + md.ell_item_off;
    const uint32_t* eitems = V.ell_items + md.ell_item_off;
    uint16_t* nl = V.nl + sbase * V.nl_width;
    const int nlw = V.nl_width;
    float4* gf = V.gf + sbase * MD2_INBOX;
    float4* xall = V.xall + sbase;
    float4* xsorted = V.xsorted + sbase;
    uint32_t* sorted = V.sorted + sbase;
    uint32_t* cell_start = V.cells + (size_t)sidx * 2 * (MDC_MAXC + 1);
    uint32_t* cursor = cell_start + MDC_MAXC + 1;
    uint16_t* rowlen = V.rowlen + sbase;
    float* bbox = V.bbox + (size_t)sidx * MDC_MAX_CL * 8;
    const float2* btype = B.btype_f + d->btype_off;
    float4* xg = B.xyzr + d->atom_off;

    uint32_t meta[K];
    int nrow[K];              // my rows' lengths
    float vx[K], vy[K], vz[K], fx[K], fy[K], fz[K];
#pragma unroll
    for (int k = 0; k < K; k++) {
        const int p = k * MD2_T + tid;
        const uint16_t a = perm[p];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        meta[k] = 3u;
        if (a != MD2_NONE) {
            q = xg[a];
            meta[k] = (uint32_t)B.kind[d->atom_off + a] | (B.envmask[d->atom_off + a] << 8);
        }
        x[p] = q;
        inbox[p] = make_uint4(0u, 0u, 0u, 0u);
        vx[k] = vy[k] = vz[k] = 0.f; fx[k] = fy[k] = fz[k] = 0.f; nrow[k] = 0;
    }
    if (tid < K * MD2_NW) { eblk[tid] = V.eblk[md.ell_desc_off + rank * (K * MD2_NW) + tid]; bact[tid] = make_int2(0, 0); }
    const float2 bt0 = btype[0];
    const float skin = d->maxrad * P.skin_factor;
    const float lim = 0.25f * skin * skin;
    const float bond_pref2 = 2.f * (float)pr.bond_prefactor;
    float ev_c = -1.f;
    int n_rebuild = 0, overflow = 0;
    bool bond_build = true;
    long long t_force = 0, t_build = 0;
    auto tick = [&]() -> long long { return PROF ? clock64() : 0ll; };
    const uint32_t xs = (uint32_t)__cvta_generic_to_shared(x);
    const uint32_t fls = (uint32_t)__cvta_generic_to_shared(inbox);
    const uint32_t cbs = (uint32_t)__cvta_generic_to_shared(cbuf);
    mdc_sync();

#define MOVES(k) ((meta[k] & 3u) == IGMK_KIND_BEAD || (meta[k] & 3u) == IGMK_KIND_CENTROID)

    // position g (rank << LOGNP | offset) -> that atom's coordinates, through distributed shared memory
    // (most partners live in this CTA -- the ordering follows the genome -- and a plain shared-memory load is cheaper)
    auto gather = [&](uint32_t g) -> float4 {
        const uint32_t la = xs + ((g & (np - 1)) << 4);
        return (g >> LOGNP) == rank ? lds_f4(la) : mdc_ld_f4(mdc_map(la, g >> LOGNP));
    };

    // deterministic sum over the CTA (warp shuffles, then the warp partials in one fixed tree)
    auto block_sum1 = [&](float v) -> float {
        float* buf = rbuf + (rb_par ^= 1) * 256;
        v = warp_sum(v);
        if (lane == 0) buf[warp] = v;
        __syncthreads();
        return warp_sum(lane < MD2_NW ? buf[lane] : 0.f);
    };
    // ... and over the cluster: every CTA drops its partial(s) into slot [rank] of every CTA, then all add the CL slots in rank order
    auto cluster_sum3 = [&](float& a, float& b, float& c) {
        a = block_sum1(a); b = block_sum1(b); c = block_sum1(c);
        const int par = (cb_par ^= 1);
        if (tid < CL) {
            const uint32_t dst = mdc_map(cbs + ((par * MDC_MAX_CL + rank) << 4), (uint32_t)tid);
            mdc_st_f32(dst, a); mdc_st_f32(dst + 4, b); mdc_st_f32(dst + 8, c);
        }
        mdc_sync();
        float sa = 0.f, sb = 0.f, sc = 0.f;
#pragma unroll
        for (int r = 0; r < CL; r++) { const float* s = cbuf + (par * MDC_MAX_CL + r) * 4; sa += s[0]; sb += s[1]; sc += s[2]; }
        a = sa; b = sb; c = sc;
    };

    auto cluster_sum1 = [&](float a) -> float {
        a = block_sum1(a);
        const int par = (cb_par ^= 1);
        if (tid < CL) mdc_st_f32(mdc_map(cbs + ((par * MDC_MAX_CL + rank) << 4), (uint32_t)tid), a);
        mdc_sync();
        float sa = 0.f;
#pragma unroll
        for (int r = 0; r < CL; r++) sa += cbuf[(par * MDC_MAX_CL + r) * 4];
        return sa;
    };

    auto set_stage = [&](double ev_scale, double env_scale) {
        ev_c = d->evfactor < 0 ? -1.f : (float)(d->evfactor * ev_scale / 3.14159265358979323846);
        if (tid < IGMK_MAX_ENVELOPES * 4) {
            const int e = tid >> 2, c = tid & 3;
            envp[tid] = e < d->n_env ? (float)(c < 3 ? d->env[e][c] * env_scale : d->env[e][3]) : 0.f;
        }
        __syncthreads();
    };

    // ---- neighbour table of the whole structure, built by the whole cluster ----
    auto rebuild = [&]() {
        const long long tb0 = tick();
        const int gt = rank * MD2_T + tid, GT = CL * MD2_T;            // my index among the cluster's threads
        float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const float4 q = x[p];
            xbx[p] = q.x; xby[p] = q.y; xbz[p] = q.z;
            xall[rank * np + p] = q;
            inbox[p] = make_uint4(0u, 0u, 0u, 0u);
            if ((meta[k] & 3u) == IGMK_KIND_BEAD) {
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
        __syncthreads();
        if (warp == 0) {
#pragma unroll
            for (int c = 0; c < 3; c++) { lo[c] = warp_min(rb[lane * 6 + c]); hi[c] = warp_max(rb[lane * 6 + 3 + c]); }
            if (lane == 0) {
#pragma unroll
                for (int c = 0; c < 3; c++) { bbox[rank * 8 + c] = lo[c]; bbox[rank * 8 + 4 + c] = hi[c]; }
            }
        }
        for (int c = gt; c < 2 * (MDC_MAXC + 1); c += GT) cell_start[c] = 0u;      // cell_start and cursor
        mdc_sync();
#pragma unroll
        for (int c = 0; c < 3; c++) { lo[c] = 1e30f; hi[c] = -1e30f; }
        for (int r = 0; r < CL; r++) {
#pragma unroll
            for (int c = 0; c < 3; c++) { lo[c] = fminf(lo[c], __ldcg(bbox + r * 8 + c)); hi[c] = fmaxf(hi[c], __ldcg(bbox + r * 8 + 4 + c)); }
        }
        const float rlist = 2.f * d->maxrad + skin, rlist2 = rlist * rlist;
        int dim[3]; float inv[3];
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float ext = hi[c] - lo[c];
            if (!(ext > 0.f)) ext = 0.f;
            int m = (int)(ext / rlist) + 1;
            if (m > MDC_MAXC_DIM) m = MDC_MAXC_DIM;
            float cw = ext / (float)m;
            if (cw < rlist) cw = rlist;
            dim[c] = m; inv[c] = 1.f / cw;
        }
        const int ncell = dim[0] * dim[1] * dim[2];
        int mycell[K];
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            mycell[k] = -1;
            if ((meta[k] & 3u) == IGMK_KIND_BEAD) {
                const float4 q = x[p];
                int cx = (int)((q.x - lo[0]) * inv[0]), cy = (int)((q.y - lo[1]) * inv[1]), cz = (int)((q.z - lo[2]) * inv[2]);
                cx = min(max(cx, 0), dim[0] - 1); cy = min(max(cy, 0), dim[1] - 1); cz = min(max(cz, 0), dim[2] - 1);
                mycell[k] = (cz * dim[1] + cy) * dim[0] + cx;
                atomicAdd(&cell_start[mycell[k] + 1], 1u);          // integer counts: independent of the order
            }
        }
        mdc_sync();
        if (rank == 0) {      // inclusive scan of cell_start[1..ncell] by one CTA (32 k cells: 32 per thread)
            const int per = (ncell + MD2_T - 1) / MD2_T;
            const int c0 = min(tid * per, ncell), c1 = min(c0 + per, ncell);
            uint32_t s = 0;
            for (int c = c0; c < c1; c++) s += __ldcg(cell_start + c + 1);
            uint32_t incl = s;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
            uint32_t* ib = (uint32_t*)(rbuf + (rb_par ^= 1) * 256);
            if (lane == 31) ib[warp] = incl;
            __syncthreads();
            uint32_t base = 0;
            for (int w = 0; w < warp; w++) base += ib[w];
            uint32_t run = base + incl - s;
            for (int c = c0; c < c1; c++) { const uint32_t cnt = __ldcg(cell_start + c + 1); run += cnt; cell_start[c + 1] = run; }
        } else {
            rb_par ^= 1;      // keep the buffer parity in step with CTA 0
        }
        mdc_sync();
#pragma unroll
        for (int k = 0; k < K; k++) {
            if (mycell[k] >= 0) {
                const uint32_t pos = atomicAdd(&cursor[mycell[k]], 1u);
                sorted[__ldcg(cell_start + mycell[k]) + pos] = (uint32_t)(rank * np + k * MD2_T + tid);
            }
        }
        mdc_sync();
        const int nbead = (int)__ldcg(cell_start + ncell);
        for (int c = gt; c < ncell; c += GT) {        // sort every (small) cell by position: the row order must not depend on the race
            const int b = (int)__ldcg(cell_start + c), e = (int)__ldcg(cell_start + c + 1);
            for (int q = b + 1; q < e; q++) {
                const uint32_t v = sorted[q]; int r = q - 1;
                while (r >= b && sorted[r] > v) { sorted[r + 1] = sorted[r]; r--; }
                sorted[r + 1] = v;
            }
        }
        mdc_sync();
        for (int s0 = gt; s0 < nbead; s0 += GT) {      // (the radius is not needed in the search: w carries the position index)
            const uint32_t a = __ldcg(sorted + s0);
            float4 v = __ldcg(xall + a);
            v.w = __uint_as_float(a);
            xsorted[s0] = v;
        }
        mdc_sync();
        // search: the cluster's threads take the beads in cell order
        int ovf = 0;
        for (int s0 = gt; s0 < nbead; s0 += GT) {
            const int p = (int)sorted[s0];
            const float4 q = xsorted[s0];
            int cx = (int)((q.x - lo[0]) * inv[0]), cy = (int)((q.y - lo[1]) * inv[1]), cz = (int)((q.z - lo[2]) * inv[2]);
            cx = min(max(cx, 0), dim[0] - 1); cy = min(max(cy, 0), dim[1] - 1); cz = min(max(cz, 0), dim[2] - 1);
            const int ax0 = max(cx - 1, 0), ax1 = min(cx + 1, dim[0] - 1);
            Md2RowPacked row(nl, p, NPS, nlw);
            for (int dz = -1; dz <= 1; dz++) {
                const int az = cz + dz;
                if (az < 0 || az >= dim[2]) continue;
                for (int dy = -1; dy <= 1; dy++) {
                    const int ay = cy + dy;
                    if (ay < 0 || ay >= dim[1]) continue;
                    const int cb = (az * dim[1] + ay) * dim[0];
                    const int b = (int)cell_start[cb + ax0], e = (int)cell_start[cb + ax1 + 1];
                    for (int s1 = b; s1 < e; s1 += 4) {        // four candidates at once, branch-free, hits packed through the table (md2.cuh)
                        float4 pj[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) pj[u] = xsorted[min(s1 + u, nbead - 1)];
                        uint32_t hits = 0u;
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            const float ddx = q.x - pj[u].x, ddy = q.y - pj[u].y, ddz = q.z - pj[u].z;
                            hits |= (uint32_t)(ddx * ddx + ddy * ddy + ddz * ddz < rlist2) << u;
                        }
                        const int rem = e - s1;
                        const uint32_t me = (uint32_t)(s0 - s1);
                        hits &= (rem >= 4 ? 0xfu : (1u << rem) - 1u) & (me < 4u ? ~(1u << me) : ~0u);
                        const uint32_t i01 = __byte_perm(__float_as_uint(pj[0].w), __float_as_uint(pj[1].w), 0x5410);
                        const uint32_t i23 = __byte_perm(__float_as_uint(pj[2].w), __float_as_uint(pj[3].w), 0x5410);
                        const uint4 L = lut[hits];
                        const uint32_t r0 = __byte_perm(i01, i23, L.x) & L.z, r1 = __byte_perm(i01, i23, L.y) & L.w;
                        row.append(((uint64_t)r1 << 32) | r0, __popc(hits));
                    }
                }
            }
            rowlen[p] = (uint16_t)row.finish((uint32_t)p, &ovf);      // padded with the atom's own index
        }
        // did any CTA overflow?  (one slot per CTA in every CTA's exchange buffer)
        ovf = __syncthreads_or(ovf);
        const int par = (cb_par ^= 1);
        if (tid < CL) mdc_st_f32(mdc_map(cbs + ((par * MDC_MAX_CL + rank) << 4), (uint32_t)tid), ovf ? 1.f : 0.f);
        mdc_sync();
#pragma unroll
        for (int r = 0; r < CL; r++) overflow |= cbuf[(par * MDC_MAX_CL + r) * 4] != 0.f;
#pragma unroll
        for (int k = 0; k < K; k++) nrow[k] = (meta[k] & 3u) == IGMK_KIND_BEAD ? (int)__ldcg(rowlen + rank * np + k * MD2_T + tid) : 0;
        if (PROF) t_build += tick() - tb0;
        bond_build = true;
        n_rebuild++;
    };

    // ---- forces on my atoms ----
    auto forces = [&](auto build_tag) {
        constexpr bool BUILD = decltype(build_tag)::value;
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const uint32_t gp = rank * np + p;
            const float4 pi = x[p];
            float ax = 0.f, ay = 0.f, az = 0.f;
            const int g = k * MD2_NW + warp;
            const int4 eb = eblk[g];
            const int2 ba = bact[g];
            {   // block A: type 0, an upper bound, inbox slot at the other end
                const uint4* it = (const uint4*)((BUILD ? eitems : eact) + eb.x) + lane;
                const int nb4 = BUILD ? eb.y : ba.x;
                uint32_t* out = eact + eb.x + (lane << 2);
                const float r0 = bt0.x, r02 = r0 * r0, kf = -bond_pref2 * bt0.y;
                const float cl = fmaxf(r0 - skin, 0.f), cl2 = cl * cl;
                int ccnt = 0;
                uint4 nxt = nb4 > 0 ? it[0] : make_uint4(0, 0, 0, 0);
                for (int c = 0; c < nb4; c++) {
                    const uint4 cur = nxt;
                    if (c + 1 < nb4) nxt = it[(c + 1) * 32];
                    const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
                    float4 pj[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) pj[u] = gather((v[u] >> 4) & 0xffffu);
                    float dx[4], dy[4], dz[4], r2[4];
                    bool any = false;
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        dx[u] = pi.x - pj[u].x; dy[u] = pi.y - pj[u].y; dz[u] = pi.z - pj[u].z;
                        r2[u] = dx[u] * dx[u] + dy[u] * dy[u] + dz[u] * dz[u];
                        any |= r2[u] > r02;
                    }
                    if (any) {
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            if (r2[u] > r02) {
                                const float fb = kf * (1.f - r0 * fast_rsqrt(r2[u]));
                                ax += dx[u] * fb; ay += dy[u] * fb; az += dz[u] * fb;
                                const uint32_t tg = (v[u] >> 4) & 0xffffu, q = v[u] & 15u;
                                gf[(size_t)tg * MD2_INBOX + q] = make_float4(dx[u], dy[u], dz[u], fb);
                                mdc_st_u8(mdc_map(fls + ((tg & (np - 1)) << 4) + q, tg >> LOGNP), 1u);
                            }
                        }
                    }
                    if (BUILD) {
#pragma unroll
                        for (int u = 0; u < 4; u++)
                            if (r2[u] > cl2) { out[((ccnt >> 2) << 7) + (ccnt & 3)] = v[u]; ccnt++; }
                    }
                }
                if (BUILD) {
                    const int most = warp_max((ccnt + 3) >> 2);
                    for (int cc = ccnt; cc < (most << 2); cc++) out[((cc >> 2) << 7) + (cc & 3)] = (gp << 4) | 15u;
                    if (lane == 0) bact[g].x = most;
                }
            }
            if (eb.w > 0) {   // block B: any type; SYM items are evaluated at both ends
                const uint4* it = (const uint4*)((BUILD ? eitems : eact) + eb.z) + lane;
                const int nb4 = BUILD ? eb.w : ba.y;
                uint32_t* out = eact + eb.z + (lane << 2);
                int ccnt = 0;
                for (int c = 0; c < nb4; c++) {
                    const uint4 cur = it[c * 32];
                    const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
                    float4 pj[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) pj[u] = gather((v[u] >> 4) & 0xffffu);
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float2 t = btype[v[u] >> MDC_IT_TSHIFT];
                        const float dx = pi.x - pj[u].x, dy = pi.y - pj[u].y, dz = pi.z - pj[u].z;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float r02 = t.x * t.x;
                        if ((t.y > 0.f ? (r2 > r02) : (r2 < r02)) && r2 > 0.f) {
                            const float fb = -bond_pref2 * fabsf(t.y) * (1.f - t.x * fast_rsqrt(r2));
                            ax += dx * fb; ay += dy * fb; az += dz * fb;
                            if (!(v[u] & MDC_IT_SYM)) {
                                const uint32_t tg = (v[u] >> 4) & 0xffffu, q = v[u] & 15u;
                                gf[(size_t)tg * MD2_INBOX + q] = make_float4(dx, dy, dz, fb);
                                mdc_st_u8(mdc_map(fls + ((tg & (np - 1)) << 4) + q, tg >> LOGNP), 1u);
                            }
                        }
                        if (BUILD) {
                            const float lo = fmaxf(t.x - skin, 0.f), hi = t.x + skin;
                            if (t.y > 0.f ? (r2 > lo * lo) : (r2 < hi * hi)) { out[((ccnt >> 2) << 7) + (ccnt & 3)] = v[u]; ccnt++; }
                        }
                    }
                }
                if (BUILD) {
                    const int most = warp_max((ccnt + 3) >> 2);
                    for (int cc = ccnt; cc < (most << 2); cc++) out[((cc >> 2) << 7) + (cc & 3)] = (gp << 4) | MDC_IT_SYM;
                    if (lane == 0) bact[g].y = most;
                }
            }
            // pair_style soft (branch-free; rows are padded with the atom's own index: r2 = 0, no force)
            if (ev_c >= 0.f) {
                const uint2* col = (const uint2*)nl + gp;
                const int nb4 = (nrow[k] + 3) >> 2;
                uint2 wn = nb4 > 0 ? col[0] : make_uint2(0u, 0u);
                for (int q = 0; q < nb4; q++) {
                    const uint2 w = wn;
                    if (q + 1 < nb4) wn = col[(size_t)(q + 1) * NPS];
                    const uint32_t j[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
                    float4 pj[4];
#pragma unroll
                    for (int u = 0; u < 4; u++) pj[u] = gather(j[u]);
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float dx = pi.x - pj[u].x, dy = pi.y - pj[u].y, dz = pi.z - pj[u].z;
                        const float rc = pi.w + pj[u].w;
                        const float r2 = dx * dx + dy * dy + dz * dz;
                        const float rinv = fast_rsqrt(r2);
                        const float sn = fast_sin(3.14159265358979323846f * (r2 * rinv) * fast_rcp(rc));
                        float fp = ev_c * rc * sn * rinv;
                        fp = (r2 < rc * rc && r2 > 0.f) ? fp : 0.f;
                        ax += dx * fp; ay += dy * fp; az += dz * fp;
                    }
                }
            }
            // fix ellipsoidalenvelope
            uint32_t m = meta[k] >> 8;
            while (m) {
                const int e = __ffs(m) - 1; m &= m - 1;
                const float ia = fast_rcp(envp[4 * e] - pi.w), ib = fast_rcp(envp[4 * e + 1] - pi.w),
                            ic = fast_rcp(envp[4 * e + 2] - pi.w), kk = envp[4 * e + 3];
                const float ux = pi.x * ia, uy = pi.y * ib, uz = pi.z * ic;
                const float k2 = ux * ux + uy * uy + uz * uz;
                const bool act = (kk > 0.f ? (k2 > 1.f) : (kk < 0.f && k2 < 1.f)) && k2 > 0.f;
                if (act) {
                    const float g2 = fabsf(kk) * (1.f - fast_rsqrt(k2));
                    ax -= g2 * pi.x; ay -= g2 * pi.y; az -= g2 * pi.z;
                }
            }
            fx[k] = ax; fy[k] = ay; fz[k] = az;
        }
        bond_build = false;
        // Every inbox flag and mailbox of this evaluation is written: oriented bonds never cross CTAs (the host lists a bond
        // between two CTAs at both ends, SYM), so this is a CTA barrier, not a cluster barrier.  The partner positions other
        // CTAs still gather from this one stay untouched until the cluster barrier of the kinetic-energy sum (step 0: below).
        __syncthreads();
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const uint4 fl = inbox[p];
            if (fl.x | fl.y | fl.z | fl.w) {
                // one bit per slot; the flagged mailboxes are fetched four at a time in slot order (as in k_md2)
                uint32_t m = (((fl.x * 0x01020408u) >> 24) & 0xfu) | ((((fl.y * 0x01020408u) >> 24) & 0xfu) << 4) |
                             ((((fl.z * 0x01020408u) >> 24) & 0xfu) << 8) | ((((fl.w * 0x01020408u) >> 24) & 0xfu) << 12);
                const float4* box = gf + ((size_t)(rank * np + p) << 4);
                while (m) {
                    float4 f[4];
#pragma unroll
                    for (int b = 0; b < 4; b++) {
                        f[b] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (m) { f[b] = __ldcg(box + (__ffs(m) - 1)); m &= m - 1; }
                    }
#pragma unroll
                    for (int b = 0; b < 4; b++) { fx[k] -= f[b].x * f[b].w; fy[k] -= f[b].y * f[b].w; fz[k] -= f[b].z * f[b].w; }
                }
                inbox[p] = make_uint4(0u, 0u, 0u, 0u);
            }
            if (!MOVES(k)) { fx[k] = 0.f; fy[k] = 0.f; fz[k] = 0.f; }
        }
    };

    // ------------------------------------------------------------------------------------
    rebuild();
    const float dt = (float)pr.timestep, dtf = 0.5f * dt;
    const float dof = 3.f * (float)d->n_nonfixed - 3.f;
    float last_temp = 0.f;
    int dbg_pass = 0;

    for (int st = (P.debug_stage >= 0 ? P.debug_stage : 0); st < pr.n_stage; st++) {
        set_stage(pr.evfactor[st], pr.envfactor[st]);
        for (int sub = (pr.relax_mdsteps > 0 ? 0 : 1); sub < 2; sub++) {
            const int nsteps = sub == 0 ? pr.relax_mdsteps : pr.mdsteps[st];
            const float t0 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstart[st];
            const float t1 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstop[st];
            const float xmax = sub == 0 ? (float)pr.relax_max_velocity : (float)pr.max_velocity;
            const float vlim2 = (xmax / dt) * (xmax / dt);
            {   // velocity nonfixed create t0 seed+st
                uint32_t seed = (uint32_t)((d->seed + st) % 2147483647ll);
                if (seed == 0) seed = 1;
                float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    float a = 0.f, b = 0.f, c = 0.f;
                    if (MOVES(k)) {
                        uint32_t r = ranpark_jump(seed, 3u * (uint32_t)perm[k * MD2_T + tid] + 1u);
                        const double AM = 1.0 / 2147483647.0;
                        a = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); b = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); c = (float)(AM * r - 0.5);
                    }
                    vx[k] = a; vy[k] = b; vz[k] = c;
                    s0 += a; s1 += b; s2 += c;
                }
                cluster_sum3(s0, s1, s2);
                const float inv = d->n_nonfixed > 0 ? 1.f / (float)d->n_nonfixed : 0.f;
                float ke = 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) {
                    if (MOVES(k)) {
                        vx[k] -= s0 * inv; vy[k] -= s1 * inv; vz[k] -= s2 * inv;
                        ke += vx[k] * vx[k] + vy[k] * vy[k] + vz[k] * vz[k];
                    }
                }
                ke = cluster_sum1(ke);
                const float tcur = dof > 0.f ? ke / dof : 0.f;
                const float fac = tcur > 0.f ? sqrtf(t0 / tcur) : 0.f;
#pragma unroll
                for (int k = 0; k < K; k++) { vx[k] *= fac; vy[k] *= fac; vz[k] *= fac; }
            }
            float scale = 1.f;
            for (int step = 0; step <= nsteps; step++) {
                const long long tc0 = tick();
                if (bond_build) forces(std::true_type{}); else forces(std::false_type{});
                if (PROF) t_force += tick() - tc0;
                if (P.debug_stage >= 0) {
                    if (dbg_pass++ == 0) { step = -1; mdc_sync(); continue; }
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        const uint16_t a = perm[k * MD2_T + tid];
                        if (a != MD2_NONE && V.dbg_forces) {
                            double* o = V.dbg_forces + 3 * (size_t)(d->atom_off + a);
                            o[0] = fx[k]; o[1] = fy[k]; o[2] = fz[k];
                        }
                    }
                    if (tid == 0 && rank == 0) B.flags[sidx] = overflow ? 1 : 0;
                    mdc_sync();
                    return;
                }
                float need_f = 0.f;
                if (step == 0) mdc_sync();      // no kinetic-energy sum in the run setup: every CTA has read the positions before any drifts
                if (step > 0) {
                    float ke = 0.f;
#pragma unroll
                    for (int k = 0; k < K; k++) {
                        float a = vx[k] + dtf * fx[k], b = vy[k] + dtf * fy[k], c = vz[k] + dtf * fz[k];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[k] = a; vy[k] = b; vz[k] = c;
                        ke += a * a + b * b + c * c;
                    }
                    ke = cluster_sum1(ke);
                    const float tcur = dof > 0.f ? ke / dof : 0.f;
                    const float ttar = t0 + ((float)step / (float)nsteps) * (t1 - t0);
                    scale = 1.f;
                    if (tcur > 0.f && fabsf(tcur - ttar) > 0.1f) scale = sqrtf(ttar / tcur);
                    last_temp = tcur * scale * scale;
                }
                if (step == nsteps) {
#pragma unroll
                    for (int k = 0; k < K; k++) { vx[k] *= scale; vy[k] *= scale; vz[k] *= scale; }
                    break;
                }
#pragma unroll
                for (int k = 0; k < K; k++) {
                    if (MOVES(k)) {
                        const int p = k * MD2_T + tid;
                        float a = vx[k] * scale + dtf * fx[k], b = vy[k] * scale + dtf * fy[k], c = vz[k] * scale + dtf * fz[k];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[k] = a; vy[k] = b; vz[k] = c;
                        float4 q = x[p];
                        q.x += dt * a; q.y += dt * b; q.z += dt * c;
                        x[p] = q;
                        const float ddx = q.x - xbx[p], ddy = q.y - xby[p], ddz = q.z - xbz[p];
                        if (ddx * ddx + ddy * ddy + ddz * ddz > lim) need_f = 1.f;
                    }
                }
                // positions of every CTA written + has any bead of the structure moved half the skin?
                need_f = cluster_sum1(need_f);
                if (need_f > 0.f) rebuild();
            }
        }
    }
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
    float badf = cluster_sum1(bad ? 1.f : 0.f);
    if (tid == 0 && rank == 0) {
        igmk_result_t* r = B.res + sidx;
        const double dof_all = 3.0 * n - 3.0;
        r->temperature = dof_all > 0 ? (double)last_temp * (double)dof / dof_all : 0.0;
        r->n_rebuilds = n_rebuild;
        B.flags[sidx] = (overflow ? 1 : 0) | (badf > 0.f ? 2 : 0);
        if (PROF && V.prof) { long long* o = V.prof + 16 * (size_t)sidx; o[0] = t_force; o[2] = t_build; o[3] = n_rebuild; }
    }
#undef MOVES
}

}  // namespace igmk
