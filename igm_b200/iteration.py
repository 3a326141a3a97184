"""One Hi-C A/M iteration on the GPU(s): what ``ActivationDistanceStep`` followed by ``ModelingStep``
do for a population (bin/igm-run:99-131), minus the file system.

    coordinates (this rank's structures, host)                      igm/steps/...
      -> upload, all-gather of the slab on the structure axis       (workers read the .hss)
      -> activation distances of this rank's chunk of the pairs     ActivationDistanceStep.task   :192-222
      -> all-gather of the rows (= actdist.hdf5)                    ActivationDistanceStep.reduce :226-290
      -> contact assignment for this rank's structures              inter/intraHiC._apply  inter_hic.py:49-64
      -> structures = template + slab + assigned bonds              ModelingStep.task             :184-264
      -> annealing MD + CG, chunk by chunk behind the assembly      Model.optimize -> lammps.optimize
      -> optimised coordinates, per-structure info, violation stats ModelingStep.task :341-382

Single process: ranks = 1.  Several GPUs: one process per GPU under torch.distributed (NCCL); every
rank passes its own contiguous block of structures (``parallel.structure_shard``).
"""
import time

import numpy as np

from . import _abi, parallel, restraints_pop
from .astep import Population
from .kernel import info_from_result, run_population


def hic_iteration(crd_local, genome, radii, template, pairs, optimizer_options, device=0, contact_range=2.0,
                  k_hic=1.0, seeds=None, bond_prefactor=1.0, violations=False, tol=0.05, astep="gather", assemble="device"):
    """crd_local: float32 [n_bead][S_local][3] (host).  genome: object with chrom, copy_off, copy_beads.
    template: PackedSystem of the shared particles and restraints (Steric, Envelope, Polymer).
    pairs: (pair_i, pair_j, pwish, plast) haploid pairs kept at this sigma -- identical on every rank.
    astep: "gather" = all-gather the coordinate slab, shard the pair list (every rank sees the whole population);
           "reduce" = the population stays sharded by structure, per-pair counts and radix histograms are all-reduced
           (``parallel.sharded_actdist``; BASELINE configs[4]).  Both give the same rows, bit for bit.
    assemble: "device" = the assignment bit matrix stays on the GPU and the M-step batches are assembled there
           (igmk_submit_population_dev); "host" = bit matrix to the host, igmk_submit_population.  Same batches, byte for byte.
    -> dict(coordinates [n_bead][S_local][3], infos [S_local], rows (row, col, dist, prob), timings,
            violations = per-structure (n_imposed, n_violated) of the harmonic restraints if asked)."""
    import torch
    rank, ws = parallel.world()
    n_bead, S = crd_local.shape[0], crd_local.shape[1]
    t0 = time.perf_counter()
    dev = "cuda:%d" % device
    local_t = torch.from_numpy(np.ascontiguousarray(crd_local, dtype=np.float32)).to(dev)
    pi, pj, pw, pl = pairs
    if astep == "gather":
        slab_t = parallel.all_gather_slab(local_t)
        with Population(slab_t.shape[1], radii, genome.chrom, genome.copy_off, genome.copy_beads, device=device) as popg:
            popg.attach(slab_t)
            rows = parallel.distributed_actdist(lambda a, b, c, d: popg.actdist(a, b, c, d, contact_range), pi, pj, pw, pl,
                                                genome.chrom, genome.copy_off, device=dev)
        del slab_t
    elif astep != "reduce":
        raise ValueError("astep must be 'gather' or 'reduce'")
    with Population(S, radii, genome.chrom, genome.copy_off, genome.copy_beads, device=device) as popl:
        popl.attach(local_t)
        if astep == "reduce":
            rows = parallel.sharded_actdist(popl, parallel.total_structures(S, device=dev), pi, pj, pw, pl, contact_range)
        row, col, dist, prob = rows
        if assemble not in ("device", "host"):
            raise ValueError("assemble must be 'device' or 'host'")
        popl.set_rows(row, col, dist)
        popl.assign_launch()
        bits = popl.assign_collect() if assemble == "host" else None
        t_astep = time.perf_counter() - t0
        r0 = (contact_range * (radii[row] + radii[col])).astype(np.float64)      # float32 product like inter_hic.py:55
        protocol = _abi.make_protocol(optimizer_options, bond_prefactor=bond_prefactor)
        out = {"rows": rows, "timings": {"astep_s": t_astep}}
        new, results, viol, tim = run_population(template, crd_local, n_bead, row, col, r0, k_hic, bits, protocol, device=device,
                                                 seeds=seeds, violations_tol=tol if violations else None,
                                                 pop=popl if assemble == "device" else None)
    infos = []
    for k, res in enumerate(results):
        if res.status != 0 or not np.isfinite(res.final_energy):
            raise RuntimeError("b200 kernel produced non-finite coordinates or energy (structure %d of rank %d)" % (k, rank))
        infos.append(info_from_result(res))
    if violations:
        out["violations"] = viol
    md_ms, cg_ms, h2d = tim["md_kernel_ms"], tim["cg_kernel_ms"], tim["h2d_bytes"]
    out["coordinates"], out["infos"] = new, infos
    out["timings"].update({"mstep_s": time.perf_counter() - t0 - t_astep, "md_kernel_ms": md_ms, "cg_kernel_ms": cg_ms,
                           "h2d_bytes": h2d, "total_s": time.perf_counter() - t0,
                           "chunks_assembled_on_device": tim.get("assembled_on_device", 0)})
    return out


def mstep_population(crd, genome, radii, template, row, col, dist, protocol, device=0, contact_range=2.0, k_hic=1.0,
                     seeds=None, violations_tol=None):
    """Assignment + M-step of one block of structures from HOST buffers with the hand-off on the device: coordinates
    (float32 [n_bead][S][3]) and the actdist rows (row, col, dist) go up once, the contact assignment
    (inter_hic.py:49-64), the assembly of every structure (ModelingStep.py:184-264) and the optimisation run on the
    GPU, the optimised coordinates come back.  -> (coordinates, [Result], violations or None, timings dict)"""
    crd = np.ascontiguousarray(crd, dtype=np.float32)
    n_bead, S = crd.shape[0], crd.shape[1]
    with Population(S, radii, genome.chrom, genome.copy_off, genome.copy_beads, device=device) as pop:
        pop.upload(crd)
        pop.set_rows(row, col, dist)
        pop.assign_launch()
        r0 = (contact_range * (radii[row] + radii[col])).astype(np.float64)
        new, results, viol, tim = run_population(template, crd, n_bead, row, col, r0, k_hic, None, protocol, device=device,
                                                 seeds=seeds, violations_tol=violations_tol, pop=pop)
    tim["h2d_bytes"] += crd.nbytes + 3 * 4 * int(np.size(row)) + 4 * int(np.size(row))     # slab, rows, bound keys
    return new, results, viol, tim


def assemble_restraints(crd_local, genome, radii, hic=None, damid=None, fish=None, device=0):
    """The A-steps of one iteration for this rank's structures: everything ``ModelingStep.task`` later reads from the
    actdist / assignment files, as arrays.

    hic:   dict(pairs=(pair_i, pair_j, pwish, plast), contact_range=2.0)                     ActivationDistanceStep
    damid: dict(loci, pwish, plast, contact_range=0.05, shape="sphere"|"ellipsoid", nucleus_param=R | (a, b, c))
                                                                                             DamidActivationDistanceStep
    fish:  dict(rtype, tol, center, probes=, radial_min=, radial_max=, pairs=, pair_min=, pair_max=) with the SORTED
           experimental distributions [n_items][S_total] of the input file (FishAssignmentStep.py:220-252); `center` =
           atom index of the restraint's static centre particle in the template                FishAssignmentStep
    -> dict(hic_rows (row, col, dist, prob), hic_bits, damid_rows (bead, dist, prob), damid_bits, damid_semiaxes,
            fish_targets {name: f8 [n][S_local]}, fish_bonds (CSR of this rank's structures), seconds)"""
    import torch
    rank, ws = parallel.world()
    t0 = time.perf_counter()
    crd_local = np.ascontiguousarray(crd_local, dtype=np.float32)
    S = crd_local.shape[1]
    dev = "cuda:%d" % device
    local_t = torch.from_numpy(crd_local).to(dev)
    slab_t, offs = parallel.all_gather_slab(local_t, return_offsets=True)
    s0 = offs[rank]                                # this rank's block in the gathered slab
    out = {}
    with Population(slab_t.shape[1], radii, genome.chrom, genome.copy_off, genome.copy_beads, device=device) as popg:
        popg.attach(slab_t)
        if hic is not None:
            pi, pj, pw, pl = hic["pairs"]
            cr = hic.get("contact_range", 2.0)
            out["hic_rows"] = parallel.distributed_actdist(lambda a, b, c, d: popg.actdist(a, b, c, d, cr), pi, pj, pw, pl,
                                                           genome.chrom, genome.copy_off, device=dev)
        if damid is not None:                      # a few thousand loci at most: every rank computes all of them
            out["damid_rows"] = popg.damid_actdist(damid["loci"], damid["pwish"], damid["plast"], damid.get("contact_range", 0.05),
                                                   damid.get("shape", "sphere"), damid["nucleus_param"])
        if fish is not None:
            tg = {}
            if fish.get("probes") is not None and len(fish["probes"]):
                _, _, rmin, rmax = popg.fish_ranks(fish["probes"])
                if "radial_min" in fish:
                    tg["radial_min"] = restraints_pop.fish_targets(rmin, fish["radial_min"])[:, s0:s0 + S]
                if "radial_max" in fish:
                    tg["radial_max"] = restraints_pop.fish_targets(rmax, fish["radial_max"])[:, s0:s0 + S]
            if fish.get("pairs") is not None and len(fish["pairs"]):
                pr = np.asarray(fish["pairs"])
                _, _, pmin, pmax = popg.fish_ranks(pr[:, 0], pr[:, 1])
                if "pair_min" in fish:
                    tg["pair_min"] = restraints_pop.fish_targets(pmin, fish["pair_min"])[:, s0:s0 + S]
                if "pair_max" in fish:
                    tg["pair_max"] = restraints_pop.fish_targets(pmax, fish["pair_max"])[:, s0:s0 + S]
            out["fish_targets"] = tg
    with Population(S, radii, genome.chrom, genome.copy_off, genome.copy_beads, device=device) as popl:
        popl.attach(local_t)
        if hic is not None:
            row, col, dist, _ = out["hic_rows"]
            out["hic_bits"] = popl.assign(row, col, dist)
        if damid is not None:
            bead, dist, _ = out["damid_rows"]
            param = np.atleast_1d(np.asarray(damid["nucleus_param"], dtype=np.float64))
            semi = np.repeat(param, 3) if param.size == 1 else param          # Damid.__init__: a = b = c = R for a sphere (damid.py:96-99)
            out["damid_semiaxes"] = semi
            out["damid_bits"] = popl.damid_assign(bead, dist, damid.get("contact_range", 0.05), semi)
    if fish is not None:
        tg = out["fish_targets"]
        out["fish_bonds"] = restraints_pop.fish_bonds(
            crd_local, genome.copy_off, genome.copy_beads, int(fish["center"]), fish["rtype"], float(fish["tol"]), float(fish.get("k", 1.0)),
            probes=fish.get("probes"), radial_min=tg.get("radial_min"), radial_max=tg.get("radial_max"),
            pairs=fish.get("pairs"), pair_min=tg.get("pair_min"), pair_max=tg.get("pair_max"))
    out["seconds"] = time.perf_counter() - t0
    return out


def population_inputs(asm, radii, n_struct, hic=None, damid=None):
    """assemble_restraints output -> the arguments of ``kernel.run_population`` / ``MStepBatch.submit_population``:
    (row, col, row_r0, row_k, bits, extras)."""
    if hic is not None:
        row, col, _, _ = asm["hic_rows"]
        cr = hic.get("contact_range", 2.0)
        r0 = (cr * (radii[row] + radii[col])).astype(np.float64)             # float32 product like inter_hic.py:55
        bits, k_hic = asm["hic_bits"], float(hic.get("k", 1.0))
    else:
        row = col = np.zeros(0, dtype=np.int32)
        r0, bits, k_hic = np.zeros(0), np.zeros((0, 1), dtype=np.uint32), 1.0
    envs = []
    if damid is not None:
        envs.append(restraints_pop.damid_envelope(asm["damid_rows"][0], asm["damid_bits"], asm["damid_semiaxes"],
                                                  damid.get("contact_range", 0.05), damid.get("k", 1.0)))
    extras = None
    if envs or "fish_bonds" in asm:
        extras = restraints_pop.extras(n_struct, bonds=asm.get("fish_bonds"), envelopes=envs)
    return row, col, r0, k_hic, bits, extras


def am_iteration(crd_local, genome, radii, template, optimizer_options, hic=None, damid=None, fish=None, device=0, seeds=None,
                 bond_prefactor=1.0, violations=False, tol=0.05):
    """One A/M iteration with any of the Hi-C, DamID and FISH restraints (bin/igm-run:99-131 for the restraint set of
    igm-config_all.json): the A-steps of ``assemble_restraints``, then the M-step of the whole block of structures with
    the per-structure restraints ModelingStep.task would add (ModelingStep.py:249-330).  `template`: PackedSystem of the
    shared particles (beads, the Envelope / Damid / Fish centre dummies) and restraints (Steric, Envelope, Polymer).
    -> dict(coordinates, infos, restraints (the assemble_restraints dict), timings, violations)."""
    rank, _ = parallel.world()
    t0 = time.perf_counter()
    crd_local = np.ascontiguousarray(crd_local, dtype=np.float32)
    n_bead, S = crd_local.shape[0], crd_local.shape[1]
    asm = assemble_restraints(crd_local, genome, radii, hic=hic, damid=damid, fish=fish, device=device)
    row, col, r0, k_hic, bits, extras = population_inputs(asm, radii, S, hic=hic, damid=damid)
    protocol = _abi.make_protocol(optimizer_options, bond_prefactor=bond_prefactor)
    new, results, viol, tim = run_population(template, crd_local, n_bead, row, col, r0, k_hic, bits, protocol, device=device,
                                             seeds=seeds, violations_tol=tol if violations else None, extras=extras)
    infos = []
    for k, res in enumerate(results):
        if res.status != 0 or not np.isfinite(res.final_energy):
            raise RuntimeError("b200 kernel produced non-finite coordinates or energy (structure %d of rank %d)" % (k, rank))
        infos.append(info_from_result(res))
    total = time.perf_counter() - t0
    return {"coordinates": new, "infos": infos, "restraints": asm, "violations": viol,
            "timings": dict(tim, astep_s=asm["seconds"], mstep_s=total - asm["seconds"], total_s=total)}


def relax_init(crd_local, template, optimizer_options, device=0, seeds=None, bond_prefactor=1.0):
    """What ``RelaxInit.task`` does for every structure of a freshly generated population
    (igm/steps/RelaxInit.py:86-178: Steric + Envelope + Polymer only, then ``model.optimize``), for this
    rank's block of structures at once.  crd_local: float32 [n_bead][S_local][3] (``generate_territories``
    output, RandomInit.py:69-102); template: PackedSystem with those three restraints.
    -> (relaxed coordinates, same layout; [info dict] * S_local)."""
    rank, _ = parallel.world()
    crd_local = np.ascontiguousarray(crd_local, dtype=np.float32)
    n_bead = crd_local.shape[0]
    none_i = np.zeros(0, dtype=np.int32)
    protocol = _abi.make_protocol(optimizer_options, bond_prefactor=bond_prefactor)
    new, results, _, _ = run_population(template, crd_local, n_bead, none_i, none_i, np.zeros(0), 1.0,
                                        np.zeros((0, 1), dtype=np.uint32), protocol, device=device, seeds=seeds)
    infos = []
    for k, res in enumerate(results):
        if res.status != 0 or not np.isfinite(res.final_energy):
            raise RuntimeError("b200 kernel produced non-finite coordinates or energy (structure %d of rank %d)" % (k, rank))
        infos.append(info_from_result(res))
    return new, infos


def hic_iteration_files(hss_in, hcs_path, hss_out, actdist_out, optimizer_options, intra_sigma, inter_sigma,
                        last_actdist=None, contact_range=2.0, nucleus_radius=None, semiaxes=None, evfactor=1.0,
                        k_hic=1.0, k_poly=1.0, k_env=1.0, device=0, seeds=None, tol=0.05, bond_prefactor=1.0,
                        plast_override=None):
    """One Hi-C A/M iteration file to file, single process: population .hss + contact matrix .hcs (+ the actdist
    file of the previous iteration) in; actdist.hdf5 and the new population .hss out -- what
    ``ActivationDistanceStep`` (setup :106-189, task :192-222, reduce :226-290) followed by ``ModelingStep``
    (task :151-391, set_structure :429-501, teardown_poller :503-527) leave on disk, with the restraint set of
    ``ModelingStep.task`` for a Hi-C run: Steric, Envelope (sphere of `nucleus_radius`, default the file's
    ``envelope/params``; or ellipsoid `semiaxes`), Polymer, inter/intra Hi-C.
    -> the dict of ``hic_iteration`` plus ``summary`` (n_imposed, n_violations over envelope, polymer and Hi-C
    restraints, as ModelingStep counts them, :449-470)."""
    import json
    from . import files, synth
    pop = files.read_hss(hss_in)
    hcs = files.read_hcs(hcs_path)
    idx = pop["index"]
    if hcs["n"] != idx.n_hap:
        raise ValueError("%s has %d haploid bins, the population's index %d" % (hcs_path, hcs["n"], idx.n_hap))
    last = files.read_actdist(last_actdist) if last_actdist else None
    pairs = files.hic_pairs_from_matrix(hcs, intra_sigma, inter_sigma, last)
    if plast_override is not None:         # diagnostics (tools/run_demo.py --no-correction): the same plast for every pair
        pairs = pairs[:3] + (np.full(pairs[0].size, float(plast_override)),)
    if nucleus_radius is None and semiaxes is None:
        nucleus_radius = float(np.asarray(pop["tree"]["envelope"]["params"]).ravel()[0])
    rad = pop["radii"]
    template = synth.pack_structure(np.zeros((idx.n_bead, 3), dtype=np.float32), rad, idx, nucleus_radius=nucleus_radius or 0.0,
                                    semiaxes=semiaxes, contact_range=contact_range, evfactor=evfactor, k_poly=k_poly, k_env=k_env,
                                    seed=0, struct_id=0)
    out = hic_iteration(pop["coordinates"], idx, rad, template, pairs, optimizer_options, device=device,
                        contact_range=contact_range, k_hic=k_hic, seeds=seeds, bond_prefactor=bond_prefactor,
                        violations=True, tol=tol)
    files.write_actdist(actdist_out, *out["rows"])
    v = np.array(out["violations"], dtype=np.int64).reshape(-1, 4)
    n_poly = int(template.bond_i.size)
    hic_imp = int(v[:, 0].sum()) - n_poly * v.shape[0]
    summary = {"n_imposed": int(v[:, 0].sum() + v[:, 2].sum()), "n_violations": int(v[:, 1].sum() + v[:, 3].sum()),
               "byrestraint": {"Envelope": {"n_imposed": int(v[:, 2].sum()), "n_violations": int(v[:, 3].sum())},
                               "Polymer+HiC": {"n_imposed": int(v[:, 0].sum()), "n_violations": int(v[:, 1].sum()),
                                               "n_imposed_polymer": n_poly * v.shape[0], "n_imposed_hic": hic_imp}}}
    n_imp, n_vio = summary["n_imposed"], summary["n_violations"]
    files.write_hss(hss_out, pop["tree"], coordinates=out["coordinates"], summary=json.dumps(summary),
                    violation=(n_vio / n_imp) if n_imp else 0.0)            # ModelingStep.py:509-519
    out["summary"] = summary
    return out
