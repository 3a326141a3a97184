// md_big.cuh -- annealing MD for structures that do not fit one SM's shared memory / registers
// (more than 8192 atoms: the 200 kb genomes of BASELINE config 4, ~30 k beads).
//
// Same algorithm, arithmetic, list layouts and results as k_md2 (md2.cuh); the difference is where
// the state lives: positions, build coordinates, velocities, forces and per-atom flags are in
// global memory (L1/L2), only the cell grid of the neighbour rebuild is in shared memory, and the
// per-thread slots are walked by a run-time loop.  One CTA per structure.  This is the
// correctness-first path for large genomes; a thread-block-cluster version that keeps the positions
// in distributed shared memory is the planned replacement.
#pragma once
#include "md2.cuh"

namespace igmk {

constexpr int BIG_CELL_DIM = 24;
constexpr int BIG_MAX_CELLS = BIG_CELL_DIM * BIG_CELL_DIM * BIG_CELL_DIM;
constexpr int BIG_FLOATS_PER_POS = 4 + 3 + 3 + 3 + 1 + 1;   // x(4) xb(3) v(3) f(3) meta(1) ncnt(1)

__host__ __device__ inline size_t md_big_smem_bytes(int K)
{
    const size_t np = (size_t)K * MD2_T;
    size_t b = ((size_t)BIG_MAX_CELLS + 2) * 4;   // cell_start
    b += np * 2 * 2;                              // cell_atoms, cell_of
    b += 2 * 256 * 4 + IGMK_MAX_ENVELOPES * 4 * 4;
    b += (size_t)K * MD2_NW * 4;                   // active bond batches per warp-slot
    return (b + 255) / 256 * 256;
}

__global__ void __launch_bounds__(MD2_T, 1)
k_md_big(Batch B, Md2View V, const Md2Desc* __restrict__ mdesc, MdParams P, float* __restrict__ gstate)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const StructDesc* d = B.desc + blockIdx.x;
    const Md2Desc md = mdesc[blockIdx.x];
    const int K = md.K, np = K * MD2_T;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = d->n_atoms;

    uint32_t* cell_start = (uint32_t*)smem;
    uint16_t* cell_atoms = (uint16_t*)(cell_start + BIG_MAX_CELLS + 2);
    uint16_t* cell_of = cell_atoms + np;
    float* rbuf = (float*)(cell_of + np);
    float* envp = rbuf + 2 * 256;
    int* bact = (int*)(envp + IGMK_MAX_ENVELOPES * 4);
    int rb_par = 0;
    bool bond_build = true;

    float* base = gstate + (size_t)md.perm_off * BIG_FLOATS_PER_POS;
    float4* x = (float4*)base;
    float* xbx = base + 4 * (size_t)np; float* xby = xbx + np; float* xbz = xby + np;
    float* vx = xbz + np; float* vy = vx + np; float* vz = vy + np;
    float* fx = vz + np; float* fy = fx + np; float* fz = fy + np;
    uint32_t* meta = (uint32_t*)(fz + np);        // kind | envmask << 8
    uint32_t* ncnt = meta + np;

    const igmk_protocol_t& pr = P.prot;
    const uint16_t* perm = V.perm + md.perm_off;
    const int32_t* eoff = V.ell_off + md.ell_desc_off;
    const uint32_t* eitems = V.ell_items + md.ell_item_off;
    uint32_t* eact = V.ell_act + md.ell_item_off;
    uint16_t* nl = V.nl + (size_t)md.perm_off * V.nl_width;
    const int nlw = V.nl_width;
    const float2* btype = B.btype_f + d->btype_off;
    float4* xg = B.xyzr + d->atom_off;

    for (int p = tid; p < np; p += MD2_T) {
        const uint16_t a = perm[p];
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t m = 3u;
        if (a != MD2_NONE) { q = xg[a]; m = (uint32_t)B.kind[d->atom_off + a] | (B.envmask[d->atom_off + a] << 8); }
        x[p] = q; meta[p] = m; ncnt[p] = 0;
        vx[p] = vy[p] = vz[p] = 0.f; fx[p] = fy[p] = fz[p] = 0.f;
    }
    const float2 bt0 = btype[0], bt1 = btype[1];
    const float skin = d->maxrad * P.skin_factor;
    const float lim = 0.25f * skin * skin;
    const float bond_pref2 = 2.f * (float)pr.bond_prefactor;
    float ev_c = -1.f;
    int n_rebuild = 0, overflow = 0;
    __syncthreads();

    auto block_sum1 = [&](float v) -> float {
        float* buf = rbuf + (rb_par ^= 1) * 256;
        v = warp_sum(v);
        if (lane == 0) buf[warp] = v;
        __syncthreads();
        return warp_sum(lane < MD2_NW ? buf[lane] : 0.f);
    };
    auto mobile = [](uint32_t m) { const uint32_t kd = m & 0xffu; return kd == IGMK_KIND_BEAD || kd == IGMK_KIND_CENTROID; };

    auto set_stage = [&](double ev_scale, double env_scale) {
        ev_c = d->evfactor < 0 ? -1.f : (float)(d->evfactor * ev_scale / 3.14159265358979323846);
        if (tid < IGMK_MAX_ENVELOPES * 4) {
            const int e = tid >> 2, c = tid & 3;
            envp[tid] = e < d->n_env ? (float)(c < 3 ? d->env[e][c] * env_scale : d->env[e][3]) : 0.f;
        }
        __syncthreads();
    };

    auto rebuild = [&]() {
        float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
        for (int p = tid; p < np; p += MD2_T) {
            const float4 q = x[p];
            xbx[p] = q.x; xby[p] = q.y; xbz[p] = q.z;
            if ((meta[p] & 0xffu) == IGMK_KIND_BEAD) {
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
        for (int c = tid; c < BIG_MAX_CELLS + 2; c += MD2_T) cell_start[c] = 0;
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
            if (m > BIG_CELL_DIM) m = BIG_CELL_DIM;
            float cw = ext / (float)m;
            if (cw < rlist) cw = rlist;
            dim[c] = m; inv[c] = 1.f / cw;
        }
        const int ncell = dim[0] * dim[1] * dim[2];
        for (int p = tid; p < np; p += MD2_T) {
            int c = MD2_NONE;
            if ((meta[p] & 0xffu) == IGMK_KIND_BEAD) {
                const float4 q = x[p];
                int cx = (int)((q.x - lo[0]) * inv[0]), cy = (int)((q.y - lo[1]) * inv[1]), cz = (int)((q.z - lo[2]) * inv[2]);
                cx = min(max(cx, 0), dim[0] - 1); cy = min(max(cy, 0), dim[1] - 1); cz = min(max(cz, 0), dim[2] - 1);
                c = (cz * dim[1] + cy) * dim[0] + cx;
                atomicAdd(&cell_start[c + 1], 1u);
            }
            cell_of[p] = (uint16_t)c;
        }
        __syncthreads();
        {   // inclusive scan: afterwards cell_start[c+1] = end of cell c
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
            int bsum = 0;
            for (int w = 0; w < warp; w++) bsum += ib[w];
            int run = bsum + incl - s;
            for (int c = c0; c < c1; c++) { const int cnt = cell_start[c + 1]; run += cnt; cell_start[c + 1] = run; }
        }
        __syncthreads();
        const int nbead = cell_start[ncell];
        __syncthreads();
        // fill every cell from its end: afterwards cell_start[c+1] = START of cell c (and the end of
        // cell c is cell_start[c+2], or nbead for the last cell); then sort each cell by index so the
        // row order does not depend on the race
        for (int p = tid; p < np; p += MD2_T) {
            const int c = cell_of[p];
            if (c != MD2_NONE) cell_atoms[atomicSub(&cell_start[c + 1], 1u) - 1u] = (uint16_t)p;
        }
        __syncthreads();
        if (tid == 0) cell_start[ncell + 1] = nbead;
        __syncthreads();
        for (int c = tid; c < ncell; c += MD2_T) {
            const int b = cell_start[c + 1], e = cell_start[c + 2];
            for (int q = b + 1; q < e; q++) {
                const uint16_t v = cell_atoms[q]; int r = q - 1;
                while (r >= b && cell_atoms[r] > v) { cell_atoms[r + 1] = cell_atoms[r]; r--; }
                cell_atoms[r + 1] = v;
            }
        }
        __syncthreads();
        int ovf = 0;
        for (int s0 = tid; s0 < nbead; s0 += MD2_T) {
            const int p = cell_atoms[s0];
            const float4 q = x[p];
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
                    const int b = cell_start[cb + ax0 + 1], e = cell_start[cb + ax1 + 2];
                    for (int s1 = b; s1 < e; s1++) {
                        const int j = cell_atoms[s1];
                        const float4 pj = x[j];
                        const float ddx = q.x - pj.x, ddy = q.y - pj.y, ddz = q.z - pj.z;
                        const float rl = q.w + pj.w + skin;
                        if (s1 != s0 && ddx * ddx + ddy * ddy + ddz * ddz < rl * rl) {
                            if (cnt < nlw) row[(((size_t)(cnt >> 2) * np) << 2) + (cnt & 3)] = (uint16_t)j;
                            cnt++;
                        }
                    }
                }
            }
            if (cnt > nlw) { ovf = 1; cnt = nlw; }
            for (int cc = cnt; cc & 3; cc++) row[(((size_t)(cc >> 2) * np) << 2) + (cc & 3)] = (uint16_t)p;
            ncnt[p] = (uint32_t)cnt;
        }
        overflow |= __syncthreads_or(ovf);
        bond_build = true;        // the next force evaluation also refreshes the bond candidate list (see md2.cuh)
        n_rebuild++;
    };

    auto forces = [&]() {
        for (int k = 0; k < K; k++) {
            const int p = k * MD2_T + tid;
            const float4 pi = x[p];
            const uint32_t m0 = meta[p];
            float ax = 0.f, ay = 0.f, az = 0.f;
            if (ev_c >= 0.f) {
                const uint2* col = (const uint2*)nl + p;
                const int nb4 = ((int)ncnt[p] + 3) >> 2;
                for (int q = 0; q < nb4; q++) {
                    const uint2 w = col[(size_t)q * np];
                    const uint32_t j[4] = {w.x & 0xffffu, w.x >> 16, w.y & 0xffffu, w.y >> 16};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const float4 pj = x[j[u]];
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
            {
                const int g = k * MD2_NW + warp;
                const bool build = bond_build;
                const int bo = eoff[g], nb4 = build ? ((eoff[g + 1] - bo) >> 7) : bact[g];
                const uint4* it = (const uint4*)((build ? eitems : eact) + bo) + lane;
                uint32_t* out = eact + bo + (lane << 2);
                int ccnt = 0;
                for (int c = 0; c < nb4; c++) {
                    const uint4 cur = it[c * 32];
                    const uint32_t v[4] = {cur.x, cur.y, cur.z, cur.w};
#pragma unroll
                    for (int u = 0; u < 4; u++) {
                        const uint32_t ti = v[u] >> 16;
                        float2 t = ti == 0 ? bt0 : bt1;
                        if (ti > 1) t = btype[ti];
                        const float4 pj = x[v[u] & 0xffffu];
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
                    for (int cc = ccnt; cc < (most << 2); cc++) out[((cc >> 2) << 7) + (cc & 3)] = (uint32_t)p;
                    if (lane == 0) bact[g] = most;
                }
            }
            uint32_t m = m0 >> 8;
            while (m) {
                const int e = __ffs(m) - 1; m &= m - 1;
                const float ia = fast_rcp(envp[4 * e] - pi.w), ib = fast_rcp(envp[4 * e + 1] - pi.w),
                            ic = fast_rcp(envp[4 * e + 2] - pi.w), kk = envp[4 * e + 3];
                const float ux = pi.x * ia, uy = pi.y * ib, uz = pi.z * ic;
                const float k2 = ux * ux + uy * uy + uz * uz;
                const bool act = (kk > 0.f ? (k2 > 1.f) : (kk < 0.f && k2 < 1.f)) && k2 > 0.f;
                const float g2 = act ? fabsf(kk) * (1.f - fast_rsqrt(k2)) : 0.f;
                ax -= g2 * pi.x; ay -= g2 * pi.y; az -= g2 * pi.z;
            }
            if (!mobile(m0)) { ax = 0.f; ay = 0.f; az = 0.f; }
            fx[p] = ax; fy[p] = ay; fz[p] = az;
        }
        bond_build = false;
    };

    rebuild();
    const float dt = (float)pr.timestep, dtf = 0.5f * dt;
    const float dof = 3.f * (float)d->n_nonfixed - 3.f;
    float last_temp = 0.f;

    for (int st = (P.debug_stage >= 0 ? P.debug_stage : 0); st < pr.n_stage; st++) {
        set_stage(pr.evfactor[st], pr.envfactor[st]);
        if (P.debug_stage >= 0) {        // forces of one run-setup evaluation, for the parity test of THIS force routine:
            forces();                    // once through the full bond lists,
            __syncthreads();
            forces();                    // once through the candidates
            __syncthreads();
            for (int p = tid; p < np; p += MD2_T) {
                const uint16_t a = perm[p];
                if (a != MD2_NONE && V.dbg_forces) {
                    double* o = V.dbg_forces + 3 * (size_t)(d->atom_off + a);
                    o[0] = mobile(meta[p]) ? fx[p] : 0.f; o[1] = mobile(meta[p]) ? fy[p] : 0.f; o[2] = mobile(meta[p]) ? fz[p] : 0.f;
                }
            }
            if (tid == 0) B.flags[blockIdx.x] = overflow ? 1 : 0;
            return;
        }
        for (int sub = (pr.relax_mdsteps > 0 ? 0 : 1); sub < 2; sub++) {
            const int nsteps = sub == 0 ? pr.relax_mdsteps : pr.mdsteps[st];
            const float t0 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstart[st];
            const float t1 = sub == 0 ? (float)pr.relax_temperature : (float)pr.tstop[st];
            const float xmax = sub == 0 ? (float)pr.relax_max_velocity : (float)pr.max_velocity;
            const float vlim2 = (xmax / dt) * (xmax / dt);
            {   // velocity nonfixed create t0 seed+st (same draws per original atom id as k_md2)
                uint32_t seed = (uint32_t)((d->seed + st) % 2147483647ll);
                if (seed == 0) seed = 1;
                float s0 = 0.f, s1 = 0.f, s2 = 0.f;
                for (int p = tid; p < np; p += MD2_T) {
                    float a = 0.f, b = 0.f, c = 0.f;
                    if (mobile(meta[p])) {
                        uint32_t r = ranpark_jump(seed, 3u * (uint32_t)perm[p] + 1u);
                        const double AM = 1.0 / 2147483647.0;
                        a = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); b = (float)(AM * r - 0.5);
                        r = (uint32_t)(((uint64_t)r * 16807ull) % 2147483647ull); c = (float)(AM * r - 0.5);
                    }
                    vx[p] = a; vy[p] = b; vz[p] = c;
                    s0 += a; s1 += b; s2 += c;
                }
                s0 = block_sum1(s0); s1 = block_sum1(s1); s2 = block_sum1(s2);
                const float inv = d->n_nonfixed > 0 ? 1.f / (float)d->n_nonfixed : 0.f;
                float ke = 0.f;
                for (int p = tid; p < np; p += MD2_T) {
                    if (mobile(meta[p])) {
                        const float a = vx[p] - s0 * inv, b = vy[p] - s1 * inv, c = vz[p] - s2 * inv;
                        vx[p] = a; vy[p] = b; vz[p] = c;
                        ke += a * a + b * b + c * c;
                    }
                }
                ke = block_sum1(ke);
                const float tcur = dof > 0.f ? ke / dof : 0.f;
                const float fac = tcur > 0.f ? sqrtf(t0 / tcur) : 0.f;
                for (int p = tid; p < np; p += MD2_T) { vx[p] *= fac; vy[p] *= fac; vz[p] *= fac; }
            }
            forces();
            float scale = 1.f;
            for (int step = 0; step <= nsteps; step++) {
                if (step > 0) {
                    forces();
                    float ke = 0.f;
                    for (int p = tid; p < np; p += MD2_T) {
                        float a = vx[p] + dtf * fx[p], b = vy[p] + dtf * fy[p], c = vz[p] + dtf * fz[p];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[p] = a; vy[p] = b; vz[p] = c;
                        ke += a * a + b * b + c * c;
                    }
                    ke = block_sum1(ke);
                    const float tcur = dof > 0.f ? ke / dof : 0.f;
                    const float ttar = t0 + ((float)step / (float)nsteps) * (t1 - t0);
                    scale = 1.f;
                    if (tcur > 0.f && fabsf(tcur - ttar) > 0.1f) scale = sqrtf(ttar / tcur);
                    last_temp = tcur * scale * scale;
                } else {
                    __syncthreads();
                }
                if (step == nsteps) {
                    for (int p = tid; p < np; p += MD2_T) { vx[p] *= scale; vy[p] *= scale; vz[p] *= scale; }
                    break;
                }
                int need = 0;
                for (int p = tid; p < np; p += MD2_T) {
                    if (mobile(meta[p])) {
                        float a = vx[p] * scale + dtf * fx[p], b = vy[p] * scale + dtf * fy[p], c = vz[p] * scale + dtf * fz[p];
                        const float vsq = a * a + b * b + c * c;
                        if (vsq > vlim2) { const float sc = sqrtf(vlim2 / vsq); a *= sc; b *= sc; c *= sc; }
                        vx[p] = a; vy[p] = b; vz[p] = c;
                        float4 q = x[p];
                        q.x += dt * a; q.y += dt * b; q.z += dt * c;
                        x[p] = q;
                        const float ddx = q.x - xbx[p], ddy = q.y - xby[p], ddz = q.z - xbz[p];
                        need |= (ddx * ddx + ddy * ddy + ddz * ddz > lim);
                    }
                }
                if (__syncthreads_or(need)) rebuild();
            }
        }
    }
    int bad = 0;
    for (int p = tid; p < np; p += MD2_T) {
        const uint16_t a = perm[p];
        if (a != MD2_NONE) {
            const float4 q = x[p];
            xg[a] = q;
            bad |= !(isfinite(q.x) && isfinite(q.y) && isfinite(q.z));
        }
    }
    bad = __syncthreads_or(bad);
    if (tid == 0) {
        igmk_result_t* r = B.res + blockIdx.x;
        const double dof_all = 3.0 * n - 3.0;
        r->temperature = dof_all > 0 ? (double)last_temp * (double)dof / dof_all : 0.0;
        r->n_rebuilds = n_rebuild;
        B.flags[blockIdx.x] = (overflow ? 1 : 0) | (bad ? 2 : 0);
    }
}

}  // namespace igmk
