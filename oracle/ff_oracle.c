/*
 * ff_oracle.c -- CPU restatement (double precision, plain C) of IGM's M-step
 * arithmetic.  TEST INFRASTRUCTURE ONLY: it is imported by tests/, by
 * __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
 * legs, never by the product path (igm_b200/ fails loudly without its CUDA
 * library and has no CPU fallback).
 *
 * What it restates.  The M-step arithmetic of alberlab/igm does not live in
 * the reference tree: igm/model/kernel/lammps.py:443-448 pipes a generated
 * script into `lmp_serial` built from github.com/alberlab/lammpgen (modified
 * LAMMPS with the USER-GENOME package; README.md:45,83-94 -- cloned at HEAD,
 * NO pinned version; absent from /root/reference and from this image).  The
 * specification followed here is therefore
 *   (1) the generated input: lammps.py:63-146 (coefficients) and :149-358
 *       (commands), lammps_model.py:253-329 (atoms / bonds),
 *   (2) the published semantics of the upstream LAMMPS commands the script
 *       uses (pair_style soft, fix adapt scale yes, fix nve/limit,
 *       fix temp/rescale, velocity create, min_style cg + quadratic line
 *       search),
 *   (3) IGM's own Python mirrors of the two custom terms:
 *       EllipticEnvelope.getScore (forces.py:206-218) and
 *       HarmonicUpperBound/LowerBound (forces.py:92-185, lammps_model.py:33-100).
 *
 * PARITY PIN STATUS
 *   pinned   : soft-sphere pair energy (E_pair) and the k>0 envelope energy
 *              (f_envelope0) against the 100 per-structure values shipped in
 *              demo/demo_sample_outputs/igm-model.hss.T (tests/golden/).
 *   UNPINNED : bond energy prefactor (E = k dr^2 assumed, LAMMPS `harmonic`
 *              convention; runtime switch bond_prefactor), the k<0 envelope
 *              branch, all force vectors, integrator / thermostat / minimiser
 *              trajectories.  "parity unpinned" for those: no lammpgen binary,
 *              source or force dump exists anywhere we can reach.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../include/igmk.h"

#define PI 3.14159265358979323846

typedef struct {
    int n;
    double *x, *v, *f;            /* [n][3] */
    const igmk_system_t *sys;
    double *rad;
    int n_nonfixed;
    /* per-envelope inverse lookup is not needed on the CPU: loop over members */
    double ev_scale, env_scale, bond_pref;
    /* verlet list over beads (half list) */
    int *nl_start, *nl_idx; int nl_cap;
    double *x_build; double skin, maxrad;
    int n_rebuild;
    double e_pair, e_bond, e_env[IGMK_MAX_ENVELOPES];
    long n_eval;
} ctx_t;

/* ---- neighbour list (results do not depend on it: LAMMPS `neighbor maxrad bin`,
 *      `neigh_modify every 1 check yes`, lammps.py:216-217) ----------------- */
static void nl_build(ctx_t *c)
{
    const int n = c->n;
    const uint8_t *kind = c->sys->kind;
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
    int nb = 0;
    for (int i = 0; i < n; i++) {
        if (kind[i] != IGMK_KIND_BEAD) continue;
        nb++;
        for (int d = 0; d < 3; d++) {
            if (c->x[3*i+d] < lo[d]) lo[d] = c->x[3*i+d];
            if (c->x[3*i+d] > hi[d]) hi[d] = c->x[3*i+d];
        }
    }
    memcpy(c->x_build, c->x, sizeof(double) * 3 * n);
    c->n_rebuild++;
    c->nl_start[0] = 0;
    if (nb == 0) { for (int i = 0; i <= n; i++) c->nl_start[i] = 0; return; }
    double rlist = 2.0 * c->maxrad + c->skin;
    int dim[3];
    for (int d = 0; d < 3; d++) {
        double ext = hi[d] - lo[d];
        int m = (int)(ext / rlist) + 1;
        if (m > 64) m = 64;
        dim[d] = m;
    }
    double cw[3];
    for (int d = 0; d < 3; d++) { cw[d] = (hi[d] - lo[d]) / dim[d]; if (cw[d] < rlist) cw[d] = rlist; }
    int ncell = dim[0] * dim[1] * dim[2];
    int *head = (int*)malloc(sizeof(int) * ncell);
    int *next = (int*)malloc(sizeof(int) * n);
    int *cell = (int*)malloc(sizeof(int) * n);
    for (int k = 0; k < ncell; k++) head[k] = -1;
    for (int i = n - 1; i >= 0; i--) {
        if (kind[i] != IGMK_KIND_BEAD) { cell[i] = -1; continue; }
        int ci[3];
        for (int d = 0; d < 3; d++) {
            int q = (int)((c->x[3*i+d] - lo[d]) / cw[d]);
            if (q >= dim[d]) q = dim[d] - 1;
            if (q < 0) q = 0;
            ci[d] = q;
        }
        cell[i] = (ci[2] * dim[1] + ci[1]) * dim[0] + ci[0];
        next[i] = head[cell[i]]; head[cell[i]] = i;
    }
    int cnt = 0;
    for (int i = 0; i < n; i++) {
        c->nl_start[i] = cnt;
        if (cell[i] < 0) continue;
        int cx = cell[i] % dim[0], cy = (cell[i] / dim[0]) % dim[1], cz = cell[i] / (dim[0] * dim[1]);
        for (int dz = -1; dz <= 1; dz++) for (int dy = -1; dy <= 1; dy++) for (int dx = -1; dx <= 1; dx++) {
            int ax = cx + dx, ay = cy + dy, az = cz + dz;
            if (ax < 0 || ay < 0 || az < 0 || ax >= dim[0] || ay >= dim[1] || az >= dim[2]) continue;
            for (int j = head[(az * dim[1] + ay) * dim[0] + ax]; j >= 0; j = next[j]) {
                if (j <= i) continue;
                double ddx = c->x[3*i] - c->x[3*j], ddy = c->x[3*i+1] - c->x[3*j+1], ddz = c->x[3*i+2] - c->x[3*j+2];
                double rl = c->rad[i] + c->rad[j] + c->skin;
                if (ddx*ddx + ddy*ddy + ddz*ddz < rl * rl) {
                    if (cnt >= c->nl_cap) {
                        c->nl_cap = c->nl_cap * 2 + 1024;
                        c->nl_idx = (int*)realloc(c->nl_idx, sizeof(int) * c->nl_cap);
                    }
                    c->nl_idx[cnt++] = j;
                }
            }
        }
    }
    c->nl_start[n] = cnt;
    free(head); free(next); free(cell);
}

static int nl_stale(const ctx_t *c)
{
    double lim = 0.25 * c->skin * c->skin;
    for (int i = 0; i < c->n; i++) {
        double dx = c->x[3*i] - c->x_build[3*i], dy = c->x[3*i+1] - c->x_build[3*i+1], dz = c->x[3*i+2] - c->x_build[3*i+2];
        if (dx*dx + dy*dy + dz*dz > lim) return 1;
    }
    return 0;
}

/* ---- energy + forces at c->x (one `run 0`) --------------------------------- */
static double energy_forces(ctx_t *c)
{
    const igmk_system_t *s = c->sys;
    const int n = c->n;
    double *x = c->x, *f = c->f;
    memset(f, 0, sizeof(double) * 3 * n);
    c->n_eval++;
    if (c->n_rebuild == 0 || nl_stale(c)) nl_build(c);

    /* pair_style soft (upstream LAMMPS), coefficients from create_lammps_data
     * lammps.py:128-135: A = ((ri+rj)/pi)^2 * evfactor, rc = ri+rj, BEAD-BEAD only;
     * per-stage scale from fix adapt, lammps.py:287-290;
     * special_bonds 1 1 1 (lammps.py:179): bonded pairs are NOT excluded. */
    double ep = 0.0;
    if (s->evfactor >= 0.0) {
        for (int i = 0; i < n; i++) {
            for (int q = c->nl_start[i]; q < c->nl_start[i+1]; q++) {
                int j = c->nl_idx[q];
                double dx = x[3*i] - x[3*j], dy = x[3*i+1] - x[3*j+1], dz = x[3*i+2] - x[3*j+2];
                double rc = (double)(float)(s->radius[i] + s->radius[j]);   /* dc computed from float32 radii */
                double r2 = dx*dx + dy*dy + dz*dz;
                if (r2 >= rc * rc) continue;
                double r = sqrt(r2);
                double A = (rc / PI) * (rc / PI) * s->evfactor * c->ev_scale;
                double arg = PI * r / rc;
                ep += A * (1.0 + cos(arg));
                if (r > 0.0) {
                    double fp = A * sin(arg) * PI / rc / r;
                    f[3*i] += dx * fp; f[3*i+1] += dy * fp; f[3*i+2] += dz * fp;
                    f[3*j] -= dx * fp; f[3*j+1] -= dy * fp; f[3*j+2] -= dz * fp;
                }
            }
        }
    }
    c->e_pair = ep;

    /* bond_style harmonic_upper_bound / harmonic_lower_bound (lammpgen);
     * activity rule from lammps_model.py:33-38,68-73 and forces.py:128-132,176-180;
     * energy convention E = pref * k * (r-r0)^2 (UNPINNED, see header). */
    double eb = 0.0;
    for (int b = 0; b < s->n_bonds; b++) {
        int i = s->bond_i[b], j = s->bond_j[b];
        double dx = x[3*i] - x[3*j], dy = x[3*i+1] - x[3*j+1], dz = x[3*i+2] - x[3*j+2];
        double r = sqrt(dx*dx + dy*dy + dz*dz);
        double dr = r - s->bond_r0[b];
        int active = (s->bond_style[b] == IGMK_BOND_UPPER) ? (dr > 0.0) : (dr < 0.0);
        if (!active) continue;
        double rk = c->bond_pref * s->bond_k[b] * dr;
        eb += rk * dr;
        if (r > 0.0) {
            double fb = -2.0 * rk / r;
            f[3*i] += dx * fb; f[3*i+1] += dy * fb; f[3*i+2] += dz * fb;
            f[3*j] -= dx * fb; f[3*j+1] -= dy * fb; f[3*j+2] -= dz * fb;
        }
    }
    c->e_bond = eb;

    /* fix ellipsoidalenvelope a b c k (lammpgen); energy mirrors
     * EllipticEnvelope.getScore forces.py:206-218 (PINNED for k>0 by the demo
     * f_envelope0 values); semi-axes scaled per stage, lammps.py:298-299;
     * k<0 (DamID, damid.py:133-139) acts inside the ellipsoid (UNPINNED). */
    double etot = ep + eb;
    for (int e = 0; e < s->n_env; e++) {
        double a = s->env_abck[4*e] * c->env_scale, bb = s->env_abck[4*e+1] * c->env_scale,
               cc = s->env_abck[4*e+2] * c->env_scale, k = s->env_abck[4*e+3];
        double ee = 0.0;
        for (int q = s->env_off[e]; q < s->env_off[e+1]; q++) {
            int i = s->env_members[q];
            double rr = c->rad[i];
            double px = x[3*i], py = x[3*i+1], pz = x[3*i+2];
            double sa = a - rr, sb = bb - rr, sc = cc - rr;
            double k2 = px*px/(sa*sa) + py*py/(sb*sb) + pz*pz/(sc*sc);
            int active = (k > 0.0) ? (k2 > 1.0) : (k2 < 1.0);
            if (!active || k == 0.0) continue;
            double nrm = sqrt(px*px + py*py + pz*pz);
            if (nrm == 0.0) continue;
            double g = 1.0 - 1.0 / sqrt(k2);
            double t = g * nrm, ak = fabs(k);
            ee += 0.5 * ak * t * t;
            f[3*i] -= ak * g * px; f[3*i+1] -= ak * g * py; f[3*i+2] -= ak * g * pz;
        }
        c->e_env[e] = ee;
        etot += ee;
    }

    /* fix 1 dummy setforce 0 0 0, lammps.py:223 */
    for (int i = 0; i < n; i++)
        if (s->kind[i] == IGMK_KIND_STATIC) f[3*i] = f[3*i+1] = f[3*i+2] = 0.0;
    return etot;
}

/* ---- Park-Miller generator used by `velocity create` (upstream RanPark) ----- */
static double ranpark(int *seed)
{
    int k = *seed / 127773;
    *seed = 16807 * (*seed - k * 127773) - 2836 * k;
    if (*seed < 0) *seed += 2147483647;
    return (1.0 / 2147483647.0) * (*seed);
}

static double temperature(const ctx_t *c, int nonfixed_only)
{
    double s = 0.0; int cnt = 0;
    for (int i = 0; i < c->n; i++) {
        if (c->sys->kind[i] == IGMK_KIND_STATIC) { if (!nonfixed_only) cnt++; continue; }
        cnt++;
        s += c->v[3*i]*c->v[3*i] + c->v[3*i+1]*c->v[3*i+1] + c->v[3*i+2]*c->v[3*i+2];
    }
    double dof = 3.0 * cnt - 3.0;
    return dof > 0 ? s / dof : 0.0;
}

/* `velocity nonfixed create T seed` with defaults dist uniform, mom yes, rot no,
 * loop all (lammps.py:322-323,335): one draw triple per atom id, group atoms
 * keep theirs, COM velocity removed, scaled to T exactly. */
static void velocity_create(ctx_t *c, double T, int64_t seed64)
{
    int seed = (int)seed64;
    if (seed <= 0) seed = 1;
    double vcm[3] = {0, 0, 0}; int cnt = 0;
    for (int i = 0; i < c->n; i++) {
        double vx = ranpark(&seed) - 0.5, vy = ranpark(&seed) - 0.5, vz = ranpark(&seed) - 0.5;
        if (c->sys->kind[i] == IGMK_KIND_STATIC) { c->v[3*i] = c->v[3*i+1] = c->v[3*i+2] = 0.0; continue; }
        c->v[3*i] = vx; c->v[3*i+1] = vy; c->v[3*i+2] = vz;
        vcm[0] += vx; vcm[1] += vy; vcm[2] += vz; cnt++;
    }
    if (cnt == 0) return;
    for (int d = 0; d < 3; d++) vcm[d] /= cnt;
    for (int i = 0; i < c->n; i++) {
        if (c->sys->kind[i] == IGMK_KIND_STATIC) continue;
        for (int d = 0; d < 3; d++) c->v[3*i+d] -= vcm[d];
    }
    double t = temperature(c, 1);
    if (t == 0.0) return;
    double fac = sqrt(T / t);
    for (int i = 0; i < 3 * c->n; i++) c->v[i] *= fac;
}

/* one `run nsteps` with fix nve/limit xmax + fix temp/rescale 1 t0 t1 0.1 1
 * (lammps.py:325-332,337-349) */
static void run_md(ctx_t *c, int nsteps, double dt, double xmax, double t0, double t1)
{
    const int n = c->n;
    const double dtf = 0.5 * dt, vlim2 = (xmax / dt) * (xmax / dt);
    energy_forces(c);                       /* setup: forces at the start of the run */
    for (int step = 1; step <= nsteps; step++) {
        for (int i = 0; i < n; i++) {
            if (c->sys->kind[i] == IGMK_KIND_STATIC) continue;
            double *v = c->v + 3*i, *f = c->f + 3*i, *x = c->x + 3*i;
            v[0] += dtf * f[0]; v[1] += dtf * f[1]; v[2] += dtf * f[2];
            double vsq = v[0]*v[0] + v[1]*v[1] + v[2]*v[2];
            if (vsq > vlim2) { double sc = sqrt(vlim2 / vsq); v[0] *= sc; v[1] *= sc; v[2] *= sc; }
            x[0] += dt * v[0]; x[1] += dt * v[1]; x[2] += dt * v[2];
        }
        energy_forces(c);
        for (int i = 0; i < n; i++) {
            if (c->sys->kind[i] == IGMK_KIND_STATIC) continue;
            double *v = c->v + 3*i, *f = c->f + 3*i;
            v[0] += dtf * f[0]; v[1] += dtf * f[1]; v[2] += dtf * f[2];
            double vsq = v[0]*v[0] + v[1]*v[1] + v[2]*v[2];
            if (vsq > vlim2) { double sc = sqrt(vlim2 / vsq); v[0] *= sc; v[1] *= sc; v[2] *= sc; }
        }
        /* temp/rescale, window 0.1, fraction 1 */
        double tcur = temperature(c, 1);
        double delta = (double)step / (double)nsteps;
        double ttar = t0 + delta * (t1 - t0);
        if (tcur > 0.0 && fabs(tcur - ttar) > 0.1) {
            double fac = sqrt(ttar / tcur);
            for (int i = 0; i < 3 * n; i++) c->v[i] *= fac;
        }
    }
}

/* min_style cg + `minimize etol ftol maxiter maxeval` (lammps.py:354-355):
 * Polak-Ribiere with the upstream quadratic line search, dmax 0.1. */
static int minimize_cg(ctx_t *c, double etol, double ftol, int maxiter, int maxeval,
                       double *efinal, int *iters, int *evals)
{
    const int nv = 3 * c->n;
    const double ALPHA_MAX = 1.0, ALPHA_REDUCE = 0.5, BACKTRACK_SLOPE = 0.4,
                 QUADRATIC_TOL = 0.1, EMACH = 1.0e-8, EPS_QUAD = 1.0e-28, EPS_ENERGY = 1.0e-8,
                 DMAX = 0.1;
    double *h = (double*)malloc(sizeof(double) * nv), *g = (double*)malloc(sizeof(double) * nv),
           *x0 = (double*)malloc(sizeof(double) * nv);
    double *x = c->x, *f = c->f;
    long eval0 = c->n_eval;
    double ecur = energy_forces(c);
    int neval = 0, stop = IGMK_CG_MAXITER, iter = 0;
    for (int i = 0; i < nv; i++) h[i] = g[i] = f[i];
    double gg = 0.0; for (int i = 0; i < nv; i++) gg += f[i] * f[i];
    (void)eval0;
    for (iter = 0; iter < maxiter; iter++) {
        double eprev = ecur;
        /* ---- line search ---- */
        int fail = -1;
        {
            double eorig = ecur;
            double fdoth = 0.0, hmax = 0.0;
            for (int i = 0; i < nv; i++) { fdoth += f[i] * h[i]; if (fabs(h[i]) > hmax) hmax = fabs(h[i]); }
            if (fdoth <= 0.0) fail = IGMK_CG_DOWNHILL;
            else if (hmax == 0.0) fail = IGMK_CG_ZEROFORCE;
            else {
                double alphamax = DMAX / hmax; if (alphamax > ALPHA_MAX) alphamax = ALPHA_MAX;
                memcpy(x0, x, sizeof(double) * nv);
                double alpha = alphamax, fhprev = fdoth, alphaprev = 0.0, engprev = eorig;
                for (;;) {
                    for (int i = 0; i < nv; i++) x[i] = x0[i] + alpha * h[i];
                    ecur = energy_forces(c); neval++;
                    double fh = 0.0; for (int i = 0; i < nv; i++) fh += f[i] * h[i];
                    double delfh = fh - fhprev;
                    if (fabs(fh) < EPS_QUAD || fabs(delfh) < EPS_QUAD) {
                        memcpy(x, x0, sizeof(double) * nv); ecur = energy_forces(c); neval++;
                        fail = IGMK_CG_ZEROQUAD; break;
                    }
                    double relerr = fabs(1.0 - (0.5 * (alpha - alphaprev) * (fh + fhprev) + ecur) / engprev);
                    double alpha0 = alpha - (alpha - alphaprev) * fh / delfh;
                    if (relerr <= QUADRATIC_TOL && alpha0 > 0.0 && alpha0 < alphamax) {
                        for (int i = 0; i < nv; i++) x[i] = x0[i] + alpha0 * h[i];
                        ecur = energy_forces(c); neval++;
                        if (ecur - eorig < EMACH) { fail = -1; break; }
                    }
                    double de_ideal = -BACKTRACK_SLOPE * alpha * fdoth;
                    double de = ecur - eorig;
                    if (de <= de_ideal) { fail = -1; break; }
                    fhprev = fh; engprev = ecur; alphaprev = alpha;
                    alpha *= ALPHA_REDUCE;
                    if (alpha <= 0.0 || de_ideal >= -EMACH) {
                        memcpy(x, x0, sizeof(double) * nv); ecur = energy_forces(c); neval++;
                        fail = IGMK_CG_ZEROALPHA; break;
                    }
                }
            }
        }
        if (fail >= 0) { stop = fail; break; }
        if (neval >= maxeval) { stop = IGMK_CG_MAXEVAL; iter++; break; }
        if (fabs(ecur - eprev) < etol * 0.5 * (fabs(ecur) + fabs(eprev) + EPS_ENERGY)) { stop = IGMK_CG_ETOL; iter++; break; }
        double d0 = 0.0, d1 = 0.0;
        for (int i = 0; i < nv; i++) { d0 += f[i] * f[i]; d1 += f[i] * g[i]; }
        if (ftol > 0.0 && d0 < ftol * ftol) { stop = IGMK_CG_FTOL; iter++; break; }
        double beta = (d0 - d1) / gg; if (beta < 0.0) beta = 0.0;
        if ((iter + 2) % nv == 0) beta = 0.0;
        gg = d0;
        double gh = 0.0;
        for (int i = 0; i < nv; i++) { g[i] = f[i]; h[i] = g[i] + beta * h[i]; gh += g[i] * h[i]; }
        if (gh <= 0.0) for (int i = 0; i < nv; i++) h[i] = g[i];
    }
    *efinal = ecur; *iters = iter; *evals = neval;
    free(h); free(g); free(x0);
    return stop;
}

static void ctx_init(ctx_t *c, const igmk_system_t *s)
{
    memset(c, 0, sizeof(*c));
    c->n = s->n_atoms; c->sys = s;
    c->x = (double*)malloc(sizeof(double) * 3 * c->n);
    c->v = (double*)calloc(3 * c->n, sizeof(double));
    c->f = (double*)calloc(3 * c->n, sizeof(double));
    c->x_build = (double*)malloc(sizeof(double) * 3 * c->n);
    c->rad = (double*)malloc(sizeof(double) * c->n);
    c->nl_start = (int*)calloc(c->n + 1, sizeof(int));
    c->nl_cap = 16 * c->n + 1024; c->nl_idx = (int*)malloc(sizeof(int) * c->nl_cap);
    c->maxrad = 0.0;
    for (int i = 0; i < c->n; i++) {
        for (int d = 0; d < 3; d++) c->x[3*i+d] = (double)s->xyz[3*i+d];
        c->rad[i] = (double)s->radius[i];
        if (s->kind[i] == IGMK_KIND_BEAD && c->rad[i] > c->maxrad) c->maxrad = c->rad[i];
        if (s->kind[i] != IGMK_KIND_STATIC) c->n_nonfixed++;
    }
    c->skin = c->maxrad;              /* `neighbor maxrad bin`, lammps.py:216 */
    c->ev_scale = 1.0; c->env_scale = 1.0; c->bond_pref = 1.0;
}

static void ctx_free(ctx_t *c)
{
    free(c->x); free(c->v); free(c->f); free(c->x_build); free(c->rad); free(c->nl_start); free(c->nl_idx);
}

/* ---- exported ----------------------------------------------------------------- */

/* `run 0`: energies and forces at the given coordinates. e_out = {total, pair, bond, env0..} */
int ofr_single_point(const igmk_system_t *s, double ev_scale, double env_scale, double bond_prefactor,
                     double *f_out, double *e_out)
{
    ctx_t c; ctx_init(&c, s);
    c.ev_scale = ev_scale; c.env_scale = env_scale; c.bond_pref = bond_prefactor;
    double e = energy_forces(&c);
    if (f_out) memcpy(f_out, c.f, sizeof(double) * 3 * c.n);
    e_out[0] = e; e_out[1] = c.e_pair; e_out[2] = c.e_bond;
    for (int k = 0; k < s->n_env; k++) e_out[3 + k] = c.e_env[k];
    ctx_free(&c);
    return 0;
}

/* the whole generated script: stages (relax + main run) then CG (Appendix A.7 of SURVEY.md) */
int ofr_optimize(const igmk_system_t *s, const igmk_protocol_t *p, double *xyz_out, igmk_result_t *res)
{
    ctx_t c; ctx_init(&c, s);
    c.bond_pref = p->bond_prefactor;
    for (int st = 0; st < p->n_stage; st++) {
        c.ev_scale = p->evfactor[st]; c.env_scale = p->envfactor[st];
        if (p->relax_mdsteps > 0) {
            velocity_create(&c, p->relax_temperature, s->seed + st);
            run_md(&c, p->relax_mdsteps, p->timestep, p->relax_max_velocity, p->relax_temperature, p->relax_temperature);
        }
        velocity_create(&c, p->tstart[st], s->seed + st);
        run_md(&c, p->mdsteps[st], p->timestep, p->max_velocity, p->tstart[st], p->tstop[st]);
    }
    c.ev_scale = 1.0;                 /* unfix exVolAdapt (reset yes) restores A, lammps.py:351 */
    memset(res, 0, sizeof(*res));
    double efinal; int iters = 0, evals = 0, stop = 0;
    if (!p->skip_minimize)
        stop = minimize_cg(&c, p->etol, p->ftol, p->max_cg_iter, p->max_cg_eval, &efinal, &iters, &evals);
    else
        efinal = energy_forces(&c);
    res->final_energy = efinal; res->pair_energy = c.e_pair; res->bond_energy = c.e_bond;
    res->n_env = s->n_env;
    for (int k = 0; k < s->n_env; k++) res->env_energy[k] = c.e_env[k];
    res->temperature = temperature(&c, 0);      /* thermo temp: group all */
    res->cg_iters = iters; res->cg_evals = evals; res->cg_stop = stop; res->n_rebuilds = c.n_rebuild;
    res->status = 0;
    for (int i = 0; i < 3 * c.n; i++) { if (!isfinite(c.x[i])) res->status = 1; xyz_out[i] = c.x[i]; }
    ctx_free(&c);
    return 0;
}

/* first `n` values of the generator for the tests */
void ofr_ranpark(int seed, int n, double *out) { for (int i = 0; i < n; i++) out[i] = ranpark(&seed); }
