/*
 * igmk.h -- C ABI of the B200-native IGM optimisation kernel ("igmk").
 *
 * This is the drop-in boundary for IGM's A/M hot path.  Every entry point is
 * extern "C", takes plain pointers / sizes / PODs and returns an int status
 * (0 = ok, non-zero = failure; the text is available from igmk_last_error()).
 * Nothing throws across the ABI.  Paths cited below are relative to the
 * alberlab/igm reference tree.
 *
 * M-step (replaces igm/model/kernel/lammps.py:361 `optimize(model, cfg)`,
 * i.e. the create_lammps_data / create_lammps_script / Popen(lmp_serial) /
 * get_info_from_log / get_last_frame sequence at lammps.py:429-465):
 *     igmk_create -> igmk_submit (once per structure) -> igmk_run
 *     -> igmk_fetch (once per structure) -> igmk_clear / igmk_destroy
 *
 * A-step (replaces igm/steps/ActivationDistanceStep.py:319 `get_actdist`,
 * the row formatting/parsing at :39,:219-222,:241 and the contact-assignment
 * predicate of igm/restraints/inter_hic.py:49-64 / intra_hic.py:50-65):
 *     igmk_pop_create -> igmk_pop_upload -> igmk_pop_actdist
 *     -> igmk_pop_assign -> igmk_submit_population (M-step of the assigned population)
 * and the sibling A-steps igmk_pop_damid_actdist / igmk_pop_damid_assign / igmk_pop_fish_ranks /
 * igmk_pop_sprite_rg2 on the same population.
 */
#ifndef IGMK_H
#define IGMK_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGMK_MAX_STAGES     32
#define IGMK_MAX_ENVELOPES  8

/* Particle kinds, igm/model/particle.py:10-12 */
#define IGMK_KIND_BEAD      0   /* Particle.NORMAL        */
#define IGMK_KIND_STATIC    1   /* Particle.DUMMY_STATIC  (frozen, no sterics) */
#define IGMK_KIND_CENTROID  2   /* Particle.DUMMY_DYNAMIC (moves, no sterics)  */

/* Bond styles, igm/model/kernel/lammps_model.py:8-9 */
#define IGMK_BOND_UPPER     0   /* harmonic_upper_bound: active iff r > r0 */
#define IGMK_BOND_LOWER     1   /* harmonic_lower_bound: active iff r < r0 */

/* why the minimiser stopped (igmk_result_t.cg_stop) */
#define IGMK_CG_MAXITER   0
#define IGMK_CG_MAXEVAL   1
#define IGMK_CG_ETOL      2
#define IGMK_CG_FTOL      3
#define IGMK_CG_DOWNHILL  4
#define IGMK_CG_ZEROFORCE 5
#define IGMK_CG_ZEROALPHA 6
#define IGMK_CG_ZEROQUAD  7

/*
 * One structure = one igm.model.Model after all restraints were applied
 * (what LammpsModel.from_model, lammps_model.py:253-301, walks).
 * All pointers are HOST pointers, borrowed for the duration of the call only.
 */
typedef struct igmk_system {
    int32_t        n_atoms;      /* len(model.particles)                               */
    const float*   xyz;          /* [n_atoms][3]  Particle.pos (float32, particle.py:4) */
    const float*   radius;       /* [n_atoms]     Particle.r                            */
    const uint8_t* kind;         /* [n_atoms]     IGMK_KIND_*                           */
    int32_t        n_bonds;      /* HarmonicUpperBound / HarmonicLowerBound forces      */
    const int32_t* bond_i;       /* [n_bonds] particle index i (forces.py:113)          */
    const int32_t* bond_j;       /* [n_bonds] particle index j                          */
    const double*  bond_k;       /* [n_bonds] spring constant f.k                       */
    const double*  bond_r0;      /* [n_bonds] bound distance  f.d                       */
    const uint8_t* bond_style;   /* [n_bonds] IGMK_BOND_*                               */
    int32_t        n_env;        /* EllipticEnvelope forces (forces.py:187), <= IGMK_MAX_ENVELOPES */
    const double*  env_abck;     /* [n_env][4]  semi-axes a,b,c and k (k<0: DamID, damid.py:133) */
    const int32_t* env_off;      /* [n_env+1]   CSR offsets into env_members            */
    const int32_t* env_members;  /* particle ids of each envelope group                 */
    double         evfactor;     /* ExcludedVolume.k (lammps_model.py:282-283); <0 = no Steric restraint */
    int64_t        seed;         /* script seed, lammps.py:159-165                      */
    int64_t        struct_id;    /* model.id (error messages only)                      */
} igmk_system_t;

/*
 * The optimisation protocol = the part of create_lammps_script
 * (lammps.py:263-356) that does not depend on the structure.
 */
typedef struct igmk_protocol {
    int32_t n_stage;                         /* custom_annealing_protocol.num_steps */
    int32_t mdsteps[IGMK_MAX_STAGES];
    double  tstart[IGMK_MAX_STAGES];
    double  tstop[IGMK_MAX_STAGES];
    double  evfactor[IGMK_MAX_STAGES];       /* fix adapt ... scale yes, lammps.py:287-290 */
    double  envfactor[IGMK_MAX_STAGES];      /* envelope_factors,        lammps.py:298-299 */
    int32_t relax_mdsteps;                   /* 0: no relax sub-run (lammps.py:315-332)    */
    double  relax_temperature;
    double  relax_max_velocity;
    double  timestep;                        /* lammps.py:242 */
    double  max_velocity;                    /* nve/limit xmax, lammps.py:227 */
    double  etol, ftol;                      /* minimize, lammps.py:355 */
    int32_t max_cg_iter, max_cg_eval;
    double  bond_prefactor;                  /* E = bond_prefactor*k*(r-r0)^2; 1.0 = LAMMPS harmonic convention */
    int32_t skip_minimize;                   /* 1: stop after the last MD stage (diagnostics) */
} igmk_protocol_t;

/* What get_info_from_log (lammps_io.py:6-37) extracts from the LAMMPS log. */
typedef struct igmk_result {
    double  final_energy;                    /* 'final-energy'  */
    double  pair_energy;                     /* 'pair-energy'   */
    double  bond_energy;                     /* 'bond-energy'   */
    double  temperature;                     /* thermo 'Temp'   */
    double  env_energy[IGMK_MAX_ENVELOPES];  /* thermo 'f_envelope<j>' */
    int32_t n_env;
    int32_t cg_iters, cg_evals, cg_stop;
    int32_t n_rebuilds;                      /* neighbour-list rebuilds during MD */
    int32_t status;                          /* 0 ok, 1 non-finite coordinates/energy */
    double  md_seconds;                      /* device time of the launch that ran this structure */
} igmk_result_t;

typedef struct igmk_handle igmk_handle_t;    /* M-step batch (one device)            */
typedef struct igmk_pop    igmk_pop_t;       /* device-resident population (A-step)  */

const char* igmk_last_error(void);
int  igmk_version(void);
int  igmk_device_count(void);
/* Measurement aid (no counterpart in the reference): rates of three micro-kernels on `device`, one 1024-thread CTA
 * per SM like the MD kernel -- dependent-chain FFMA (TFLOP/s, 2 flops per FMA), the same as packed fma.rn.f32x2,
 * and conflict-free 16-byte shared-memory reads (TB/s).  bench.py states the MD kernel's FP32 / shared-memory
 * fractions against these ("confirm by micro-benchmark").  Any pointer may be NULL. */
int  igmk_peak_probe(int device, double* fp32_tflops, double* fp32x2_tflops, double* smem_tbs);

/* ---- M-step: per-structure submit / run / fetch -------------------------------- */
/* device: a CUDA ordinal, or IGMK_STAGING_ONLY for a handle that only stages on the host (submit, submit_population,
 * slot_sizes, layout_stats -- what pack + create_lammps_data do before any LAMMPS process exists, lammps.py:429-431);
 * every call that needs the GPU fails on such a handle.  There is no CPU compute path. */
#define IGMK_STAGING_ONLY (-1)
int igmk_create(int device, igmk_handle_t** out);
int igmk_destroy(igmk_handle_t* h);
int igmk_clear(igmk_handle_t* h);
/* copies the structure into the batch; *slot_out indexes it in fetch calls */
int igmk_submit(igmk_handle_t* h, const igmk_system_t* sys, int32_t* slot_out);
/*
 * Submit structures s_begin .. s_begin+n_struct-1 of a population in one call: what n_struct
 * ModelingStep.task calls assemble one by one (igm/steps/ModelingStep.py:184-264).  `tmpl` holds what
 * all structures share (particles incl. dummies, Steric / Envelope / Polymer restraints); the first
 * n_bead atoms take their coordinates from the .hss-layout slab `crd` [n_bead][n_struct_total][3]
 * (host), the others keep tmpl->xyz.  On top, structure s gets HarmonicUpperBound((row[r], col[r]),
 * row_r0[r], row_k) for every actdist row r whose bit (s - s_begin) is set in bits[r]
 * (`bits_words` words per row: the matrix igmk_pop_assign returns for a population of those
 * structures; inter_hic.py:53-62, intra_hic.py:54-63).  tmpl->struct_id + s identifies structure s; its velocity
 * seed is seeds[s - s_begin], or with seeds == NULL ((max(tmpl->seed, 1) * (tmpl->struct_id + s)) mod 9190037) + 1
 * -- the script's own formula (lammps.py:159-165) with tmpl->seed in the place of seed * step_no -- so that no two
 * structures share a `velocity create` stream.  Multi-threaded on the host.
 */
int igmk_submit_population(igmk_handle_t* h, const igmk_system_t* tmpl, int32_t n_bead,
                           int32_t n_struct_total, int32_t s_begin, int32_t n_struct, const float* crd,
                           int64_t n_rows, const int32_t* row, const int32_t* col, const double* row_r0,
                           double row_k, const uint32_t* bits, int32_t bits_words, const int64_t* seeds,
                           int32_t* first_slot_out);
/*
 * The same with restraints that differ from structure to structure beyond the Hi-C rows -- what ModelingStep.task adds
 * after inter/intraHiC (igm/steps/ModelingStep.py:268-330).  All pointers are host pointers; index k of the
 * per-structure arrays is structure s_begin + k.
 *   bonds: structure k gets bond_i/j/k/r0/style[bond_off[k] .. bond_off[k+1]) appended after its Hi-C bonds, atom
 *     indices in the template's numbering (FISH: lower / upper bounds to the restraint's static centre particle or
 *     between the selected copies, with that structure's own target distances, igm/restraints/fish.py:93-274);
 *   envelopes: n_env groups appended after the template's; group e has the parameters env_abck[e] and, in structure
 *     k, the members env_bead[r] of its rows r = env_row_off[e] .. env_row_off[e+1] whose bit k is set in
 *     env_bits[r][env_bits_words] (the matrix igmk_pop_damid_assign returns; DamID: EllipticEnvelope(affected
 *     particles of THIS structure, semiaxes * (1 - contact_range), -k), igm/restraints/damid.py:116-142).
 */
typedef struct igmk_pop_extras {
    const int64_t*  bond_off;       /* [n_struct + 1] or NULL: no extra bonds */
    const int32_t*  bond_i;
    const int32_t*  bond_j;
    const double*   bond_k;
    const double*   bond_r0;
    const uint8_t*  bond_style;     /* IGMK_BOND_* */
    int32_t         n_env;
    const double*   env_abck;       /* [n_env][4] */
    const int64_t*  env_row_off;    /* [n_env + 1] */
    const int32_t*  env_bead;
    const uint32_t* env_bits;
    int32_t         env_bits_words;
    /* particles of a structure's own, appended after the template's atoms: structure k gets atom_xyz / atom_radius /
     * atom_kind[atom_off[k] .. atom_off[k+1]) as atoms tmpl->n_atoms, tmpl->n_atoms + 1, ... (SPRITE: one DUMMY_DYNAMIC
     * centroid per cluster assigned to the structure, igm/restraints/sprite.py:52-65; its bonds come through bond_*) */
    const int64_t*  atom_off;       /* [n_struct + 1] or NULL */
    const float*    atom_xyz;       /* [..][3] */
    const float*    atom_radius;
    const uint8_t*  atom_kind;      /* IGMK_KIND_* */
} igmk_pop_extras_t;
int igmk_submit_population_ex(igmk_handle_t* h, const igmk_system_t* tmpl, int32_t n_bead,
                              int32_t n_struct_total, int32_t s_begin, int32_t n_struct, const float* crd,
                              int64_t n_rows, const int32_t* row, const int32_t* col, const double* row_r0,
                              double row_k, const uint32_t* bits, int32_t bits_words, const int64_t* seeds,
                              const igmk_pop_extras_t* extras /* or NULL */, int32_t* first_slot_out);
/* builds the MD kernel's host-side layout (atom order + sliced-ELL bond blocks) for what is submitted and reports
 * it: slots per thread, bond items (both ends), items after padding, seconds the build took (any pointer may be NULL) */
int igmk_layout_stats(igmk_handle_t* h, int32_t* slots_per_thread, int64_t* bond_items, int64_t* ell_items,
                      double* build_seconds);
/* sizes of a submitted structure (any pointer may be NULL) */
int igmk_slot_sizes(igmk_handle_t* h, int32_t slot, int32_t* n_atoms, int64_t* n_bonds, int64_t* n_env_members);
/* host -> device copy of everything submitted (what writing + re-reading the .data
 * file is in the reference, lammps.py:430-431); igmk_run does it implicitly if needed */
int igmk_upload(igmk_handle_t* h);
/* put the coordinates back to what was submitted (repeat a run on the same batch) */
int igmk_reset_coordinates(igmk_handle_t* h);
/* ONE launch sequence for everything submitted: annealing MD + CG minimisation
 * (lammps.py:285-356).  Coordinates are updated in the batch. */
int igmk_run(igmk_handle_t* h, const igmk_protocol_t* prot);
/* `run 0` equivalent: energies + forces at the submitted coordinates, double
 * precision, with the pair prefactor scaled by ev_scale and the envelope
 * semi-axes by env_scale. */
int igmk_single_point(igmk_handle_t* h, double ev_scale, double env_scale,
                      double bond_prefactor);
/* same single point in float32 through the team-per-atom routine of mstep.cuh (the arithmetic of the double-precision
 * path at MD precision; NOT the force loop of the MD kernel -- that one is igmk_md_forces) */
int igmk_single_point_f32(igmk_handle_t* h, double ev_scale, double env_scale,
                          double bond_prefactor);
/* Forces exactly as the annealing-MD kernel computes them (its float32 arithmetic, neighbour tables, oriented bond
 * lists and hardware approximations): the run-setup evaluation of annealing stage `stage` of `prot` (pair prefactor
 * and envelope scaled as lammps.py:287-303 does for that stage), no time step taken; read them with
 * igmk_fetch_forces.  What a `run 0` with the dump's fx fy fz columns is in the reference (lammps.py:253-256).
 * Available for whichever MD kernel the batch takes (one CTA, thread-block cluster, global-memory). */
int igmk_md_forces(igmk_handle_t* h, const igmk_protocol_t* prot, int32_t stage);
int igmk_fetch(igmk_handle_t* h, int32_t slot, float* xyz_out /* [n_atoms][3] */,
               igmk_result_t* res);
/* every structure of the batch at once, back into the .hss-layout slab crd [n_bead][n_struct_total][3] at
 * structures s_begin .. s_begin+n-1 (what ModelingStep.set_structure copies .hms by .hms into the master
 * matrix, `self._hss_crd[:,i,:] = crd`, igm/steps/ModelingStep.py:429-441); res: n results or NULL */
int igmk_fetch_population(igmk_handle_t* h, int32_t n_bead, int32_t n_struct_total, int32_t s_begin, float* crd,
                          igmk_result_t* res);
int igmk_fetch_forces(igmk_handle_t* h, int32_t slot, double* f_out /* [n_atoms][3] */);
/*
 * Violation ratios of the structure's restraints at the batch's CURRENT coordinates (after igmk_run:
 * the optimised ones) -- what ModelingStep.task collects force by force
 * (igm/steps/ModelingStep.py:345-372).  bond_ratio [n_bonds], submit order: HarmonicUpperBound /
 * LowerBound.getViolationRatio (igm/model/forces.py:128-136, :176-184).  env_u [sum of envelope
 * members], envelope by envelope in member order: the scale-free part of EllipticEnvelope.getScores
 * (:220-250): for k > 0, ratio = max(0, (u/scale)*k) / (k*scale); for k < 0, ratio = max(0, u*k)/(k*scale).
 * All structures are evaluated in one pass on the first call after the coordinates changed.
 */
int igmk_violation_ratios(igmk_handle_t* h, int32_t slot, double* bond_ratio, double* env_u);
/* device time (ms) of the MD and CG launches of the last igmk_run and the bytes the last
 * upload moved host -> device (any pointer may be NULL) */
int igmk_last_timings(igmk_handle_t* h, double* md_ms, double* cg_ms, int64_t* h2d_bytes);

/*
 * igmk_single_point_noalloc -- no counterpart in the reference; added for hosts that are not
 * allowed to touch the CUDA allocator: a service that grabbed one arena at start-up and runs with
 * allocation disabled afterwards, a caller inside a latency-critical loop where an implicit
 * device synchronisation from cudaMalloc/cudaFree is unacceptable, or code running where the
 * allocator may fail because another tenant filled the card.  It evaluates energies and forces of
 * ONE structure (the `run 0` of igmk_single_point, double precision) and calls no cudaMalloc,
 * cudaFree or other allocating routine: the neighbour table the kernel needs as scratch lives in a
 * fixed ~0.72 MB array inside the library image, which every GPU that loads the library keeps until
 * the process exits.  Hence the size limit: n_atoms <= IGMK_NOALLOC_MAX_ATOMS, at most 96
 * neighbours within the list cutoff per bead (status 1 otherwise); larger problems are refused with
 * an error -- use igmk_single_point for those.
 *
 * All d_* pointers are DEVICE pointers owned by the caller, in the layouts igmk_submit builds:
 *   d_xyzr [n][4] x,y,z,radius float32; d_kind [n] IGMK_KIND_*; d_envmask [n] bit e = member of
 *   envelope e; d_bond_ptr [n+1] + d_bond_items: every bond listed under both of its atoms as
 *   partner | type << 16; d_btype [n_types][2] = (r0, k) with k < 0 for a lower bound.
 *   env_abck is a HOST array [n_env][4]; max_radius = largest bead radius.
 * Outputs (device): d_forces [n][3], d_energies [3 + IGMK_MAX_ENVELOPES] = total, pair, bond,
 * envelope 0.., d_status: 0 ok, 1 neighbour table full, 2 non-finite energy.
 * The launch is asynchronous on `stream` (a cudaStream_t, NULL = default stream); read the outputs
 * after synchronising it.  The first call on each device also looks up the scratch array's address
 * and raises the kernel's shared-memory limit -- make that first call before starting any stream
 * capture.  All calls on a device share the one scratch array: nothing in it has to start at zero
 * (each call rewrites every entry it reads), but two calls running on the GPU at the same time
 * would overwrite each other's neighbour table -- keep them on one stream or order them yourself.
 * The call selects `device` for its own work and restores the caller's current device before it returns.  The
 * scratch address found on the first call belongs to the device's primary context: after a cudaDeviceReset the
 * library has to be reloaded.
 */
#define IGMK_NOALLOC_MAX_ATOMS 3328
int igmk_single_point_noalloc(int device, void* stream, int32_t n_atoms,
                              const float* d_xyzr, const uint8_t* d_kind, const uint32_t* d_envmask,
                              const int32_t* d_bond_ptr, const uint32_t* d_bond_items, const double* d_btype,
                              int32_t n_env, const double* env_abck, double evfactor, float max_radius,
                              double ev_scale, double env_scale, double bond_prefactor,
                              double* d_forces, double* d_energies, int32_t* d_status);

/* ---- A-step: device-resident population ------------------------------------------ */
/*
 * Population geometry.  `crd` uses the .hss layout coordinates[bead][struct][3]
 * float32 (ModelingStep.py:441,517).  copy_off/copy_beads are Index.copy_index
 * flattened (haploid bin -> bead ids); chrom_hap[i] = index.chrom[i] for the
 * haploid bin i (ActivationDistanceStep.py:366-370).
 */
int igmk_pop_create(int device, int32_t n_bead, int32_t n_struct, int32_t n_hap,
                    const float* radii /* [n_bead] */,
                    const int32_t* chrom_bead /* [n_bead] index.chrom */,
                    const int32_t* copy_off /* [n_hap+1] */,
                    const int32_t* copy_beads,
                    igmk_pop_t** out);
int igmk_pop_destroy(igmk_pop_t* p);
int igmk_pop_upload(igmk_pop_t* p, const float* crd /* [n_bead][n_struct][3] */);
int igmk_pop_download(igmk_pop_t* p, float* crd);
/* device pointer to the coordinate slab (for callers that keep data resident) */
int igmk_pop_coords_devptr(igmk_pop_t* p, void** dptr);
/* use a caller-owned device slab [n_bead][n_struct][3] float32 instead (e.g. a torch
 * tensor that NCCL all-gathers into); the caller keeps it alive */
int igmk_pop_attach_coords(igmk_pop_t* p, void* dptr);
/* device time (ms, CUDA events on the population's stream) of the last *_launch */
double igmk_pop_last_kernel_ms(igmk_pop_t* p);

/*
 * get_actdist (ActivationDistanceStep.py:319-459) for a batch of haploid pairs,
 * including the "%6d %6d %10.4f %.4f" text round trip and float32 parse of
 * :39/:241.  Inputs: pair_i/pair_j haploid indices, pwish/plast as float64
 * (what the .in.npy batches hold, :179).  Outputs, sized for the worst case
 * n_pairs*max_copies^2 rows: row/col bead ids, dist/prob float32 exactly as
 * they would land in actdist.hdf5.  *n_rows_out = rows produced; rows appear
 * in pair order, copy pairs in the reference's emission order.
 */
int igmk_pop_actdist(igmk_pop_t* p, int64_t n_pairs,
                     const int32_t* pair_i, const int32_t* pair_j,
                     const double* pwish, const double* plast,
                     double contact_range,
                     int64_t max_rows, int32_t* row, int32_t* col,
                     float* dist, float* prob, int64_t* n_rows_out);

/* The same in three phases, so that a caller can keep the batch resident and time the
 * kernel alone: stage the pair batch on the device, launch, copy back + expand to rows. */
int igmk_pop_set_pairs(igmk_pop_t* p, int64_t n_pairs,
                       const int32_t* pair_i, const int32_t* pair_j,
                       const double* pwish, const double* plast);
int igmk_pop_actdist_launch(igmk_pop_t* p, double contact_range);
int igmk_pop_actdist_collect(igmk_pop_t* p, int64_t max_rows, int32_t* row, int32_t* col,
                             float* dist, float* prob, int64_t* n_rows_out);

/*
 * The same activation distances with the population SHARDED by structure over several GPUs (north_star: "NCCL ...
 * reduces the per-pair distance histograms"; BASELINE configs[4]): every rank holds its own structures only.  The
 * two population-wide numbers get_actdist needs per pair -- the contact count (:419) and the o-th smallest squared
 * distance (:425, :446) -- are sums over structures:
 *     igmk_pop_set_pairs (same pair list on every rank)
 *     igmk_pop_shard_count -> d_counts [n_pairs] u32            | caller: all-reduce(sum) over the ranks
 *     igmk_pop_shard_rank  (total counts -> probability, wanted rank o; clears d_prefix)
 *     for pass = 0..3:  igmk_pop_shard_hist -> d_hist [n_pairs][256] u32   | caller: all-reduce(sum)
 *                       igmk_pop_shard_pick (the radix digit holding rank o; pass 3 also stores the distance)
 *     igmk_pop_actdist_collect (rows, identical on every rank and bit-identical to igmk_pop_actdist over the
 *     whole population)
 * d_* are DEVICE buffers owned by the caller (e.g. torch tensors handed to NCCL); each call has completed on the
 * device when it returns.
 */
int igmk_pop_shard_count(igmk_pop_t* p, double contact_range, void* d_counts);
int igmk_pop_shard_rank(igmk_pop_t* p, int32_t n_struct_total, const void* d_counts_total, void* d_want, void* d_prefix);
int igmk_pop_shard_hist(igmk_pop_t* p, int32_t pass, const void* d_prefix, void* d_hist);
int igmk_pop_shard_pick(igmk_pop_t* p, int32_t pass, const void* d_hist_total, void* d_want, void* d_prefix);

/*
 * Contact frequency of haploid pairs in the resident population: the `pnow` of get_actdist
 * (ActivationDistanceStep.py:373,:419-422), count(d_sq <= (contact_range*(ri+rj))^2) over all copy
 * pairs and structures divided by their number, as float32.  freq: host, [n_pairs].
 */
int igmk_pop_contact_freq(igmk_pop_t* p, int64_t n_pairs,
                          const int32_t* pair_i, const int32_t* pair_j,
                          double contact_range, float* freq);

/*
 * Contact assignment: for every actdist row r and every structure s decide
 * `np.linalg.norm(x_i - x_j) <= dist[r]` with the reference arithmetic
 * (inter_hic.py:53, intra_hic.py:54, particle.py:35-36).  Output is a bit
 * matrix bits[r][ceil(n_struct/32)] (bit s%32 of word s/32), host memory.
 */
int igmk_pop_assign(igmk_pop_t* p, int64_t n_rows,
                    const int32_t* row, const int32_t* col, const float* dist,
                    uint32_t* bits);
int igmk_pop_set_rows(igmk_pop_t* p, int64_t n_rows,
                      const int32_t* row, const int32_t* col, const float* dist);
int igmk_pop_assign_launch(igmk_pop_t* p);
int igmk_pop_assign_collect(igmk_pop_t* p, uint32_t* bits);

/*
 * Assignment -> M-step batch without leaving the device.  The reference turns every
 * satisfied actdist row of a structure into a HarmonicUpperBound Python object (inter_hic.py:49-64,
 * intra_hic.py:50-65) that LammpsModel re-walks (lammps_model.py:279-301); igmk_submit_population does the same
 * from a downloaded bit matrix on the host threads.  Here the slab, the rows and the bit matrix stay in the
 * igmk_pop_t the A-step used:
 *   igmk_pop_set_rows + igmk_pop_assign_launch      the bit matrix, on the device
 *   igmk_pop_set_row_bounds                          row_r0[r] = contact_range * (radius_i + radius_j) of every staged
 *                                                    row (inter_hic.py:55-62) and the common spring constant
 *   igmk_submit_population_dev(h, tmpl, pop, ...)    structures s_begin .. s_begin+n_struct-1 of the population into the
 *                                                    EMPTY batch h (same GPU): template + slab coordinates + one upper
 *                                                    bound per row whose bit is set, in row order -- five kernels
 *                                                    (assemble.cuh) write the atoms, the bond CSR, the bond type tables
 *                                                    and the MD kernel's layout.  The arrays are byte-identical to what
 *                                                    igmk_submit_population + igmk_upload build from the same inputs.
 * Seeds as in igmk_submit_population.  Returns 0, 1 (error) or 2: this input is outside what the device path covers
 * (more than 3840 atoms, more than 1024 distinct bound distances, per-structure extras) and must go through
 * igmk_submit_population; igmk_last_error says which.  Afterwards the batch behaves like any other (run, fetch,
 * violation ratios); igmk_submit cannot add to it.
 */
int igmk_pop_set_row_bounds(igmk_pop_t* p, int64_t n_rows, const double* row_r0, double row_k);
int igmk_submit_population_dev(igmk_handle_t* h, const igmk_system_t* tmpl, igmk_pop_t* pop, int32_t s_begin,
                               int32_t n_struct, const int64_t* seeds, int32_t* first_slot_out);
/*
 * Test aid: one array of the staged batch as it sits on the device (0 desc, 1 xyzr, 2 xyzr0, 3 kind, 4 envmask,
 * 5 bond_ptr, 6 bond_items, 7 btype_f, 8 btype_d, 9 md2 descriptors, 10 atom permutation, 11 block descriptors,
 * 12 sliced-ELL items).  *n_bytes = its size; copied to `out` when it fits max_bytes.
 */
int igmk_debug_fetch_array(igmk_handle_t* h, int32_t which, void* out, int64_t max_bytes, int64_t* n_bytes);

/*
 * DamID (lamina) A-step.  igmk_pop_damid_actdist = get_damid_actdist
 * (igm/steps/DamidActivationDistanceStep.py:285-356) for a batch of loci (haploid indices), with
 * the "%6d %10.2f %.5f" text round trip (:31, :229).  ellipsoid = 0: nucleus_param[0] is the
 * radius (`sphere`); 1: nucleus_param[0..2] are the semi-axes.  One output row per copy of each
 * locus, always (p = 0 gives dist 2.0, :347).  igmk_pop_damid_assign = the predicate of
 * Damid._apply_envelope (igm/restraints/damid.py:116-130) for every (row, structure):
 * snormsq_ellipsoid(pos, semiaxes*(1-contact_range), r) >= dist^2; bits as in igmk_pop_assign.
 */
int igmk_pop_damid_actdist(igmk_pop_t* p, int64_t n_loci, const int32_t* loc,
                           const double* pwish, const double* plast, double contact_range,
                           int32_t ellipsoid, const double* nucleus_param,
                           int64_t max_rows, int32_t* row_bead, float* dist, float* prob,
                           int64_t* n_rows_out);
int igmk_pop_damid_assign(igmk_pop_t* p, int64_t n_rows, const int32_t* bead, const float* dist,
                          double contact_range, const double* semiaxes, uint32_t* bits);

/*
 * FISH A-step (igm/steps/FishAssignmentStep.py:31-88): for item q = pair (hap_i, hap_j) or radial
 * probe (hap_j < 0) the minimum and maximum over the diploid copies of the pair / radial distance
 * in every structure (min_dist, max_dist: [n_items][n_struct] float32 = the values numpy holds) and
 * the rank of each structure in the population (argsort(argsort(.)), rank_*: [n_items][n_struct]);
 * the caller then gathers target[rank] like FishAssignmentStep.task (:220-225, :247-252).
 */
int igmk_pop_fish_ranks(igmk_pop_t* p, int64_t n_items, const int32_t* hap_i, const int32_t* hap_j,
                        float* min_dist, float* max_dist, int32_t* rank_min, int32_t* rank_max);

/*
 * SPRITE A-step numerics = get_rgs2 / get_rg2s_cpp (igm/cython_compiled/sprite.pyx:36-101,
 * cpp_sprite_assignment.cpp:79-143) for a batch of clusters taken straight from the resident
 * population: cluster c has regions cl_off[c]..cl_off[c+1]; region i has the alternative beads
 * beads[rg_off[i]..rg_off[i+1]] (its chromosome copies).  Outputs: rg2 [n_clusters][n_struct] the
 * smallest squared radius of gyration over all copy combinations, copy_idx [n_regions][n_struct]
 * the chosen alternative per region (int8), min_struct [n_clusters] the first structure with the
 * smallest rg2.  float32 arithmetic in the reference's order: results are bit-identical.
 */
int igmk_pop_sprite_rg2(igmk_pop_t* p, int32_t n_clusters, const int32_t* cl_off, const int32_t* rg_off,
                        const int32_t* beads, float* rg2, int8_t* copy_idx, int32_t* min_struct);

/*
 * SPRITE A-step for a batch of clusters: compute_gyration_radius (igm/cython_compiled/sprite.pyx:104-283) followed by
 * the keep-best selection of SpriteAssignmentStep.task (igm/steps/SpriteAssignmentStep.py:148-160).  Cluster c has the
 * segments cl_off[c]..cl_off[c+1] (haploid ids in ascending order, i.e. grouped by chromosome); segment i has the
 * copies beads[seg_off[i]..seg_off[i+1]) and belongs to chromosome group seg_group[i] (0.. within its cluster, at most
 * 8); rep[c][g] = the representative segment (position within the cluster) on which the copy of group g is chosen --
 * the reference draws it with np.random.choice (:222-223), here the caller does, so that a run can be replayed.
 * Outputs per cluster: the `keep` structures with the smallest squared radius of gyration in ascending order
 * (best_rg2, best_struct: [n_clusters][keep]) and, per segment, the bead selected in each of them
 * (best_beads [n_segments][keep]).  float32 arithmetic in the reference's order: bit-identical values.
 */
int igmk_pop_sprite_select(igmk_pop_t* p, int32_t n_clusters, const int32_t* cl_off, const int32_t* seg_off,
                           const int32_t* beads, const int32_t* seg_group, const int32_t* rep /* [n_clusters][8] */,
                           int32_t keep, float* best_rg2, int32_t* best_struct, int32_t* best_beads);

#ifdef __cplusplus
}
#endif
#endif /* IGMK_H */
