// assemble.cuh -- assignment -> M-step batch ON THE DEVICE.
//
// ModelingStep.task builds a Model per structure: template restraints plus one HarmonicUpperBound per actdist row the
// structure satisfies (inter_hic.py:49-64, intra_hic.py:50-65), which LammpsModel then re-walks (lammps_model.py:279-301).
// igmk_submit_population does that on the host from a downloaded bit matrix.  Here the coordinate slab, the actdist
// rows and the assignment bit matrix stay where the A-step left them (an igmk_pop_t) and five kernels, one CTA per
// structure, write the batch the MD / CG kernels read: atoms, bond CSR, bond type tables, and the oriented sliced-ELL
// layout of k_md2.  Every stage reproduces the ORDER of the host path (bonds in submit order, type ids by use count,
// the greedy owner orientation bond by bond), so the arrays are byte-identical to what igmk_submit_population +
// igmk_upload produce -- tests/test_gpu_assemble.py compares them array by array.
//
//   k_asm_count   bonds and distinct bond types per structure (sizes the host needs to place the structures)
//   k_asm_bonds   atoms; bond list in submit order (ordered compaction of the bit column); type ranks; CSR, rows
//                 sorted by bond index; position of every entry's twin in the partner's row
//   k_asm_orient  the greedy orientation: one thread walks the bonds in host order (the decision for a bond depends
//                 on the counters all earlier bonds left behind); one byte per bond comes out
//   k_asm_layout  atom order by owned items (stable rank), positions, block sizes of every warp-slot
//   k_asm_ell     the sliced-ELL items
#pragma once
#include <climits>
#include "mstep.cuh"
#include "md2.cuh"

namespace igmk {

constexpr int ASM_T = 1024;
constexpr int ASM_MAX_KEYS = 1024;        // distinct (r0, k, style) over the template and all rows
constexpr int ASM_MAX_ATOMS = 3840;       // the one-CTA MD kernel's limit (5 atoms per thread x 768 threads)

// decision byte of a bond, stored at its entry in the row of the smaller atom index a (partner b > a)
constexpr uint8_t ASM_DEC_Q = 0x0f;       // inbox slot at the receiving end
constexpr uint8_t ASM_DEC_A = 0x10;       // a lists the item
constexpr uint8_t ASM_DEC_B = 0x20;       // b lists the item
constexpr uint8_t ASM_DEC_SYM = 0x40;     // no inbox: evaluated at every end that lists it

struct AsmIn {
    int32_t n_atoms, n_bead, n_tbonds, n_keys;
    const int32_t* t_i; const int32_t* t_j; const int32_t* t_key;       // template bonds
    const float4* t_xyzr; const uint8_t* t_kind; const uint32_t* t_envmask;
    const double* key_r0; const double* key_k; const int32_t* key_style; // bond type keys (rows + template)
    const float* crd; int32_t n_struct_total, s_begin;                   // .hss-layout slab of the population
    int64_t n_rows; const int32_t* row; const int32_t* col; const int32_t* row_key;
    const uint32_t* bits; int32_t bits_words;                            // bit s of row r: structure s satisfies row r
};

struct AsmOut {
    const StructDesc* desc;
    float4* xyzr; float4* xyzr0; uint8_t* kind; uint32_t* envmask;
    int32_t* bond_ptr; uint32_t* bond_items; float2* btype_f; double2* btype_d;
    int32_t* bl_i; int32_t* bl_j; int32_t* bl_key;     // [sum nb] bonds in submit order (kept for the violation report)
    int32_t* bidx;        // [2 sum nb] bond index of every CSR entry
    int32_t* twin;        // [2 sum nb] CSR position of the same bond in the partner's row (-1: bond to oneself)
    uint8_t* dec;         // [2 sum nb] ASM_DEC_* at the entry of the smaller end
    uint16_t* cntA; uint16_t* cntB; uint16_t* pos;     // [atoms] items owned in block A / B; position in the MD ordering
    uint16_t* perm; int4* eblk; int32_t* ell_size;     // layout
    const Md2Desc* mdesc; uint32_t* ell_items;
    int32_t K;
};

__global__ void __launch_bounds__(ASM_T)
k_asm_count(AsmIn I, int32_t* __restrict__ nb_out, int32_t* __restrict__ ntypes_out)
{
    __shared__ int use[ASM_MAX_KEYS];
    __shared__ int red[2][32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sg = I.s_begin + blockIdx.x;
    for (int g = tid; g < I.n_keys; g += ASM_T) use[g] = 0;
    __syncthreads();
    for (int b = tid; b < I.n_tbonds; b += ASM_T) use[I.t_key[b]] = 1;
    const uint32_t* w = I.bits + (sg >> 5);
    const uint32_t m = 1u << (sg & 31);
    int cnt = 0;
    for (int64_t r = tid; r < I.n_rows; r += ASM_T)
        if (w[(size_t)r * I.bits_words] & m) { cnt++; use[I.row_key[r]] = 1; }     // (same value from every writer)
    __syncthreads();
    int nt = 0;
    for (int g = tid; g < I.n_keys; g += ASM_T) nt += use[g];
    cnt = warp_sum(cnt); nt = warp_sum(nt);
    if (lane == 0) { red[0][warp] = cnt; red[1][warp] = nt; }
    __syncthreads();
    if (warp == 0) {
        cnt = warp_sum(red[0][lane]); nt = warp_sum(red[1][lane]);
        if (lane == 0) { nb_out[blockIdx.x] = I.n_tbonds + cnt; ntypes_out[blockIdx.x] = nt; }
    }
}

// dynamic shared memory: use / first / tnew [n_keys], ptr [n_atoms + 1], cur [n_atoms]
__host__ __device__ inline size_t asm_bonds_smem(int n_keys, int n_atoms) { return ((size_t)3 * n_keys + 2 * (size_t)n_atoms + 1) * sizeof(int); }

__global__ void __launch_bounds__(ASM_T)
k_asm_bonds(AsmIn I, AsmOut O)
{
    extern __shared__ int asm_sm[];
    __shared__ int wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int s = blockIdx.x, sg = I.s_begin + s;
    const StructDesc d = O.desc[s];
    const int n = I.n_atoms, G = I.n_keys, nt = I.n_tbonds;
    int* use = asm_sm; int* first = use + G; int* tnew = first + G; int* ptr = tnew + G; int* cur = ptr + n + 1;
    const int64_t bl0 = d.bond_item_off / 2;             // bond_item_off = 2 * (bonds of the structures before)
    int32_t* bl_i = O.bl_i + bl0; int32_t* bl_j = O.bl_j + bl0; int32_t* bl_key = O.bl_key + bl0;
    uint32_t* items = O.bond_items + d.bond_item_off;
    int32_t* bidx = O.bidx + d.bond_item_off;
    int32_t* twin = O.twin + d.bond_item_off;

    // ---- atoms: beads from the slab, the rest from the template (ModelingStep.py:184-192) ----
    for (int a = tid; a < n; a += ASM_T) {
        float4 q = I.t_xyzr[a];
        if (a < I.n_bead) {
            const float* c = I.crd + ((size_t)a * I.n_struct_total + sg) * 3;
            q.x = c[0]; q.y = c[1]; q.z = c[2];
        }
        O.xyzr[d.atom_off + a] = q; O.xyzr0[d.atom_off + a] = q;
        O.kind[d.atom_off + a] = I.t_kind[a]; O.envmask[d.atom_off + a] = I.t_envmask[a];
    }
    for (int g = tid; g < G; g += ASM_T) { use[g] = 0; first[g] = INT_MAX; }
    for (int a = tid; a <= n; a += ASM_T) ptr[a] = 0;
    for (int a = tid; a < n; a += ASM_T) cur[a] = 0;
    __syncthreads();

    // ---- bond list in submit order: template bonds, then the rows whose bit is set, in row order ----
    for (int b = tid; b < nt; b += ASM_T) {
        const int i = I.t_i[b], j = I.t_j[b], key = I.t_key[b];
        bl_i[b] = i; bl_j[b] = j; bl_key[b] = key;
        atomicAdd(&ptr[i + 1], 1); atomicAdd(&ptr[j + 1], 1);
        atomicAdd(&use[key], 1); atomicMin(&first[key], b);
    }
    const uint32_t* w = I.bits + (sg >> 5);
    const uint32_t m = 1u << (sg & 31);
    int run = nt;
    for (int64_t base = 0; base < I.n_rows; base += ASM_T) {
        const int64_t r = base + tid;
        const bool flag = r < I.n_rows && (w[(size_t)r * I.bits_words] & m);
        const unsigned bal = __ballot_sync(0xffffffffu, flag);
        if (lane == 0) wsum[warp] = __popc(bal);
        __syncthreads();
        int incl = wsum[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        const int excl = __shfl_sync(0xffffffffu, incl, warp) - wsum[warp];
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        if (flag) {
            const int b = run + excl + __popc(bal & ((1u << lane) - 1u));
            const int i = I.row[r], j = I.col[r], key = I.row_key[r];
            bl_i[b] = i; bl_j[b] = j; bl_key[b] = key;
            atomicAdd(&ptr[i + 1], 1); atomicAdd(&ptr[j + 1], 1);
            const unsigned grp = __match_any_sync(bal, key);          // nearly every row has the same key: one atomic per warp
            if (lane == __ffs(grp) - 1) { atomicAdd(&use[key], __popc(grp)); atomicMin(&first[key], b); }
        }
        run += total;
        __syncthreads();
    }
    const int nb = run;

    __syncthreads();
    // ---- bond types: ids by use count, ties by first appearance (igmk_submit: stable sort of first-seen order) ----
    int ntypes = 0;
    for (int g = tid; g < G; g += ASM_T) {
        int rank = -1;
        const int u = use[g], f = first[g];
        if (u > 0) {
            rank = 0;
            for (int q = 0; q < G; q++) { const int uq = use[q]; rank += (uq > u) || (uq == u && first[q] < f); }
            const double r0 = I.key_r0[g], ks = I.key_style[g] == IGMK_BOND_LOWER ? -I.key_k[g] : I.key_k[g];
            O.btype_f[d.btype_off + rank] = make_float2((float)r0, (float)ks);
            O.btype_d[d.btype_off + rank] = make_double2(r0, ks);
        }
        tnew[g] = rank;
    }
    __syncthreads();
    {
        int c = 0;
        for (int g = tid; g < G; g += ASM_T) c += use[g] > 0;
        c = warp_sum(c);
        if (lane == 0) wsum[warp] = c;
        __syncthreads();
        ntypes = 0;
        for (int q = 0; q < 32; q++) ntypes += wsum[q];
        __syncthreads();
    }
    if (tid < 2 && tid >= ntypes) {                     // the kernels read entries 0 and 1 unconditionally
        O.btype_f[d.btype_off + tid] = make_float2(0.f, 0.f);
        O.btype_d[d.btype_off + tid] = make_double2(0.0, 0.0);
    }

    // ---- CSR: inclusive scan of the degrees held in ptr[1..n] ----
    {
        const int per = (n + ASM_T - 1) / ASM_T;
        const int a0 = min(tid * per, n), a1 = min(a0 + per, n);
        int sum = 0;
        for (int a = a0; a < a1; a++) sum += ptr[a + 1];
        int incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        if (lane == 31) wsum[warp] = incl;
        __syncthreads();
        int basew = 0;
        for (int q = 0; q < warp; q++) basew += wsum[q];
        int acc = basew + incl - sum;
        for (int a = a0; a < a1; a++) { acc += ptr[a + 1]; ptr[a + 1] = acc; }
        __syncthreads();
    }
    for (int a = tid; a <= n; a += ASM_T) O.bond_ptr[d.bond_ptr_off + a] = ptr[a];
    // entries in any order, then every row sorted by bond index = the order igmk_submit appends them
    for (int b = tid; b < nb; b += ASM_T) {
        const int i = bl_i[b], j = bl_j[b];
        const uint32_t t = (uint32_t)tnew[bl_key[b]] << 16;
        const int c1 = ptr[i] + atomicAdd(&cur[i], 1);
        items[c1] = (uint32_t)j | t; bidx[c1] = b;
        const int c2 = ptr[j] + atomicAdd(&cur[j], 1);
        items[c2] = (uint32_t)i | t; bidx[c2] = b;
    }
    __syncthreads();
    for (int a = tid; a < n; a += ASM_T) {
        const int b0 = ptr[a], b1 = ptr[a + 1];
        for (int q = b0 + 1; q < b1; q++) {
            const int kb = bidx[q]; const uint32_t kv = items[q];
            int r = q - 1;
            while (r >= b0 && bidx[r] > kb) { bidx[r + 1] = bidx[r]; items[r + 1] = items[r]; r--; }
            bidx[r + 1] = kb; items[r + 1] = kv;
        }
    }
    __syncthreads();
    // twin of every entry: the entry of the same bond in the partner's row (rows are sorted by bond index)
    for (int a = tid; a < n; a += ASM_T) {
        for (int c = ptr[a]; c < ptr[a + 1]; c++) {
            const int o = (int)(items[c] & 0xffffu), key = bidx[c];
            int tw = -1;
            if (o != a) {
                int lo = ptr[o], hi = ptr[o + 1] - 1;
                while (lo < hi) { const int mid = (lo + hi) >> 1; if (bidx[mid] < key) lo = mid + 1; else hi = mid; }
                tw = lo;
            }
            twin[c] = tw;
        }
    }
}

// dynamic shared memory: own, indeg, cA, cB as uint16 [n_atoms]
__host__ __device__ inline size_t asm_orient_smem(int n_atoms) { return (size_t)4 * n_atoms * sizeof(uint16_t); }

// The orientation rule of md2_build_owner (igmk.cu), bond by bond in its order: a ascending, the row of a in CSR order,
// partners b > a.  Serial by nature (every decision reads the counters the earlier ones wrote): lane 0 of one warp per
// structure; the structures of a batch run side by side.
__global__ void __launch_bounds__(32)
k_asm_orient(AsmIn I, AsmOut O)
{
    extern __shared__ uint16_t asm_su[];
    const int lane = threadIdx.x;
    const StructDesc d = O.desc[blockIdx.x];
    const int n = I.n_atoms;
    uint16_t* own = asm_su; uint16_t* indeg = own + n; uint16_t* cA = indeg + n; uint16_t* cB = cA + n;
    for (int a = lane; a < n; a += 32) { own[a] = 0; indeg[a] = 0; cA[a] = 0; cB[a] = 0; }
    __syncwarp();
    if (lane == 0) {
        const int32_t* bp = O.bond_ptr + d.bond_ptr_off;
        const uint32_t* items = O.bond_items + d.bond_item_off;
        uint8_t* dec = O.dec + d.bond_item_off;
        const uint8_t* kind = I.t_kind;
        const bool a_ok = O.btype_f[d.btype_off].y > 0.f;
        for (int a = 0; a < n; a++) {
            const int c0 = bp[a], c1 = bp[a + 1], da = c1 - c0;
            const bool fa = kind[a] == IGMK_KIND_STATIC;
            for (int c = c0; c < c1; c++) {
                const uint32_t it = items[c];
                const int b = (int)(it & 0xffffu);
                if (b <= a) continue;
                const uint32_t t = it >> 16;
                const bool fb = kind[b] == IGMK_KIND_STATIC;
                uint8_t dd;
                if (fa || fb) {                                   // frozen end: the bond lives at the moving end, without inbox
                    dd = ASM_DEC_SYM;
                    if (!fa) { dd |= ASM_DEC_A; own[a]++; cB[a]++; }
                    if (!fb) { dd |= ASM_DEC_B; own[b]++; cB[b]++; }
                } else {
                    const int db = bp[b + 1] - bp[b];
                    int o = (own[a] < own[b] || (own[a] == own[b] && da >= db)) ? a : b, r = o == a ? b : a;
                    if (indeg[r] >= MD2_INBOX - 1) { const int x = o; o = r; r = x; }
                    if (indeg[r] >= MD2_INBOX - 1) {
                        dd = ASM_DEC_SYM | ASM_DEC_A | ASM_DEC_B;
                        own[a]++; own[b]++; cB[a]++; cB[b]++;
                    } else {
                        dd = (uint8_t)((o == a ? ASM_DEC_A : ASM_DEC_B) | indeg[r]);
                        indeg[r]++; own[o]++;
                        if (a_ok && t == 0) cA[o]++; else cB[o]++;
                    }
                }
                dec[c] = dd;
            }
        }
    }
    __syncwarp();
    for (int a = lane; a < n; a += 32) { O.cntA[d.atom_off + a] = cA[a]; O.cntB[d.atom_off + a] = cB[a]; }
}

// dynamic shared memory: own uint16 [n_atoms], cols int [2 * K * 32]
__host__ __device__ inline size_t asm_layout_smem(int n_atoms, int K) { return ((size_t)n_atoms + 1) / 2 * 4 + (size_t)2 * K * MD2S_NW * sizeof(int); }

__global__ void __launch_bounds__(ASM_T)
k_asm_layout(AsmIn I, AsmOut O)
{
    extern __shared__ int asm_sl[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int s = blockIdx.x, n = I.n_atoms, K = O.K, np = K * MD2S_T, G = K * MD2S_NW;
    const StructDesc d = O.desc[s];
    uint16_t* own = (uint16_t*)asm_sl;
    int* cols = asm_sl + (n + 1) / 2;                    // [G][2]
    uint16_t* perm = O.perm + (size_t)s * np;
    const uint16_t* cA = O.cntA + d.atom_off; const uint16_t* cB = O.cntB + d.atom_off;
    uint16_t* pos = O.pos + d.atom_off;
    for (int a = tid; a < n; a += ASM_T) own[a] = (uint16_t)(cA[a] + cB[a]);
    for (int p = tid; p < np; p += ASM_T) perm[p] = MD2_NONE;
    __syncthreads();
    // stable sort by owned items, descending: rank = atoms that own more + atoms before me that own as many
    for (int a = tid; a < n; a += ASM_T) {
        const int my = own[a];
        int q = 0;
        for (int x = 0; x < n; x++) { const int o = own[x]; q += (o > my) || (o == my && x < a); }
        const int g = q >> 5, l = q & 31;
        const int k = g / MD2S_NW, r = g % MD2S_NW;
        const int gi = k * MD2S_NW + ((k & 1) ? MD2S_NW - 1 - r : r);     // boustrophedon over the warps of the MD kernel
        const int p = gi * 32 + l;
        perm[p] = (uint16_t)a; pos[a] = (uint16_t)p;
    }
    __syncthreads();
    for (int p = tid; p < np; p += ASM_T) {              // a warp of this kernel = 32 consecutive positions = one warp-slot
        const uint16_t a = perm[p];
        int ca = 0, cb = 0;
        if (a != MD2_NONE) { ca = cA[a]; cb = cB[a]; }
        ca = warp_max(ca); cb = warp_max(cb);
        if (lane == 0) { cols[2 * (p >> 5)] = (ca + 3) / 4 * 4; cols[2 * (p >> 5) + 1] = (cb + 3) / 4 * 4; }
    }
    __syncthreads();
    if (tid == 0) {
        int4* eb = O.eblk + (size_t)s * G;
        int at = 0;
        for (int gi = 0; gi < G; gi++) {
            const int c0 = cols[2 * gi], c1 = cols[2 * gi + 1];
            const int off0 = at; at += c0 * 32;
            const int off1 = at; at += c1 * 32;
            eb[gi] = make_int4(off0, c0 / 4, off1, c1 / 4);
        }
        O.ell_size[s] = at;
    }
}

__global__ void __launch_bounds__(ASM_T)
k_asm_ell(AsmIn I, AsmOut O)
{
    const int tid = threadIdx.x, lane = tid & 31;
    const int s = blockIdx.x, K = O.K, np = K * MD2S_T, G = K * MD2S_NW;
    const StructDesc d = O.desc[s];
    const Md2Desc md = O.mdesc[s];
    const uint16_t* perm = O.perm + (size_t)s * np;
    const int4* eblk = O.eblk + (size_t)s * G;
    uint32_t* out = O.ell_items + md.ell_item_off;
    const int32_t* bp = O.bond_ptr + d.bond_ptr_off;
    const uint32_t* items = O.bond_items + d.bond_item_off;
    const int32_t* twin = O.twin + d.bond_item_off;
    const uint8_t* dec = O.dec + d.bond_item_off;
    const uint16_t* pos = O.pos + d.atom_off;
    const bool a_ok = O.btype_f[d.btype_off].y > 0.f;
    for (int p = tid; p < np; p += ASM_T) {
        const int4 eb = eblk[p >> 5];
        // padding = an item to oneself; layout [batch of 4 columns][lane][4]
        for (int c = 0; c < eb.y * 4; c++) out[eb.x + ((c >> 2) * 32 + lane) * 4 + (c & 3)] = ((uint32_t)p << 4) | 15u;
        for (int c = 0; c < eb.w * 4; c++) out[eb.z + ((c >> 2) * 32 + lane) * 4 + (c & 3)] = ((uint32_t)p << 4) | MD2_IT_SYM;
        const uint16_t a16 = perm[p];
        if (a16 == MD2_NONE) continue;
        const int a = a16, c0 = bp[a], c1 = bp[a + 1];
        int nA = 0, nB = 0;
        auto emit = [&](int c, uint8_t dd) {
            const uint32_t it = items[c];
            const uint32_t t = it >> 16;
            const bool sym = dd & ASM_DEC_SYM;
            const uint32_t v = ((uint32_t)pos[it & 0xffffu] << 4) | (sym ? MD2_IT_SYM : (uint32_t)(dd & ASM_DEC_Q)) | (t << MD2_IT_TSHIFT);
            if (a_ok && t == 0 && !sym) { out[eb.x + ((nA >> 2) * 32 + lane) * 4 + (nA & 3)] = v; nA++; }
            else { out[eb.z + ((nB >> 2) * 32 + lane) * 4 + (nB & 3)] = v; nB++; }
        };
        // items pushed while a smaller atom walked its row: in the order of that walk = by twin position
        int last = -1;
        for (;;) {
            int best = INT_MAX, bc = -1;
            for (int c = c0; c < c1; c++) {
                if ((int)(items[c] & 0xffffu) >= a) continue;
                const int tw = twin[c];
                if (tw > last && tw < best && (dec[tw] & ASM_DEC_B)) { best = tw; bc = c; }
            }
            if (bc < 0) break;
            emit(bc, dec[best]);
            last = best;
        }
        // then the items of my own walk
        for (int c = c0; c < c1; c++) {
            if ((int)(items[c] & 0xffffu) <= a) continue;
            const uint8_t dd = dec[c];
            if (dd & ASM_DEC_A) emit(c, dd);
        }
    }
}

// SubBond records of a device-assembled batch, for k_bond_violations (M10)
__global__ void __launch_bounds__(256)
k_asm_sub_bonds(int64_t n_bonds, const int32_t* __restrict__ bl_i, const int32_t* __restrict__ bl_j, const int32_t* __restrict__ bl_key,
                const double* __restrict__ key_r0, const double* __restrict__ key_k, const int32_t* __restrict__ key_style,
                const int64_t* __restrict__ bond_start, int32_t n_struct, const StructDesc* __restrict__ desc,
                SubBond* __restrict__ out, int32_t* __restrict__ atom_off_of_bond)
{
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < n_bonds; b += (int64_t)gridDim.x * blockDim.x) {
        int lo = 0, hi = n_struct - 1;                   // structure of bond b: last s with bond_start[s] <= b
        while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (bond_start[mid] <= b) lo = mid; else hi = mid - 1; }
        const int key = bl_key[b];
        SubBond q;
        q.i = bl_i[b]; q.j = bl_j[b]; q.r0 = key_r0[key]; q.k = key_k[key]; q.style = key_style[key];
        q.f32 = ((double)(float)q.r0 == q.r0) ? 1 : 0;
        out[b] = q;
        atom_off_of_bond[b] = desc[lo].atom_off;
    }
}

}  // namespace igmk
