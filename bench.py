#!/usr/bin/env python
"""bench.py -- NNI candidates scored/s (and CLV site-updates/s) on N B200s.

Workload (BASELINE.json configs[2]/[3] shape, the one the north_star target is
quoted on): synthetic 64 taxa x 1,000,000 sites, GTR+Gamma4 (DS1's model).
One STEP = one pass of the hot path over one parent tree:
    P matrices of all 2n-3 branches            (kernel 2)
    the directional CLVs the candidates read   (kernel 1, one launch; 105 of the
                                                3(n-2) = 186: none that faces a tip)
    every NNI candidate of Wanderer::QueueMoves (2(n-2) = 124) scored with
    SingleBranchOptimizer semantics            (prepare + cooperative Newton)
`value`  = candidates/s with the alignment resident in HBM, CUDA events on the
           engine's stream, max over ranks.
`e2e`    = the same metric through the C ABI starting from the raw alignment in
           pinned HOST memory: the step additionally uploads and encodes the n
           tip sequences (H2D inside the timed region) and reads the results back.
N > 1: candidate-parallel: alignment and model replicated, every
rank scores the candidates of its own parent tree, no data-path collective
(weak scaling); only the timing max goes through NCCL.
`--impl reference`: the UNMODIFIED reference (oracle/_ref/libpllref.so, built in
place from /root/reference by oracle/Makefile.ref) timed on the host cores with
the identical call sequence of SingleBranchOptimizer::EvaluateMove, one private
partition per thread, on a bounded site sample, scaled to the full site count.
`--workload ptw-ds1`: BASELINE configs[1] on DS1 instead -- the whole topography walk
(`--lnl-offset -5 --optimization-radius 1`, one wanderer) through ptw_batched (pt's own
driver objects + this engine + the batched move tester) next to the reference's
ptw_threads on the host cores; candidates/s = EvaluateMove questions / wall time.
`--workload sharded`: BASELINE configs[2] -- one tree's full-tree lnL + OptimizeAllBranches with the
SITES sharded over the GPUs (strong scaling; allreduce of per-tree scalars only), see run_sharded.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from pt_b200.pytree import Tree, cand_array, random_tree_newick  # noqa: E402
from pt_b200.synth import (FREQS, RATES, SEED_ALN, SEED_TREE, SUBST, parent_tree,  # noqa: E402
                           simulate)


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons during the timed region.

    Sampled in-process through NVML (nvidia_ml_py): two light queries every 25 ms.  A polling
    `nvidia-smi -lms` child (the recipe's clocks line) is only the fallback: its NVML attach
    and its polls take driver locks that were seen to stall the benchmark's own launches for
    tens of ms inside the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        super().__init__(daemon=True)
        self.device, self.rows, self._stop_evt = device, [], threading.Event()
        self.source = None

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [x.strip() for x in vis.split(",") if x.strip()]
            if self.device < len(ids) and ids[self.device].isdigit():
                return int(ids[self.device])
        return self.device

    def _run_nvml(self):
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self._physical_index())
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        bits = (("hw_slowdown", nv.nvmlClocksThrottleReasonHwSlowdown),
                ("hw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                ("sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                ("sw_power_cap", nv.nvmlClocksThrottleReasonSwPowerCap))
        self.source = "nvml"
        while not self._stop_evt.is_set():
            sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            why = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            active = ["Active" if why & b else "Not Active" for _, b in bits]
            self.rows.append([time.time(), str(self.device), str(sm), str(mx), "", ""] + active)
            self._stop_evt.wait(0.025)

    def run(self):
        try:
            self._run_nvml()
            return
        except Exception:
            self.source = "nvidia-smi"
        try:
            p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                  "-lms", "200", "-i", str(self.device)], stdout=subprocess.PIPE,
                                 stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        while not self._stop_evt.is_set():
            line = p.stdout.readline()
            if not line:
                break
            self.rows.append([time.time()] + [x.strip() for x in line.split(",")])
        p.terminate()

    def stop(self, t0=None, t1=None):
        """summary of the samples taken inside [t0, t1] (the timed region); the sampler itself
        is started before the warm-up so that its start-up never falls inside the timed region"""
        self._stop_evt.set()
        rows = [r[1:] for r in self.rows if t0 is None or t0 <= r[0] <= t1]
        if not rows:
            rows = [r[1:] for r in self.rows]
        sm = [float(r[1]) for r in rows if len(r) > 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if len(r) > 8 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            if len(r) > 8:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                    "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None,
                    sm_max_mhz=max(mx) if mx else None, reasons=sorted(reasons),
                    samples=len(sm), source=self.source)


def kernel_sources_stamp():
    import hashlib
    h = hashlib.sha256()
    for f in ("kernels.cuh", "blo.cuh", "engine.cu"):
        with open(os.path.join(ROOT, "pt_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_traffic(kernel, taxa, sites):
    """dram__bytes_read + dram__bytes_write per launch of `kernel` from the committed
    `ncu --set full` capture (profiles/r2_traffic.json): valid for the default workload only, and
    only while the kernel sources are the ones the capture was taken on (the file carries their
    hash) -- otherwise None, not a stale number"""
    if (taxa, sites) != (64, 1000000):
        return None
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            t = json.load(f)
        if t.get("kernel_sources_sha16") != kernel_sources_stamp():
            return None
        return t[kernel]["dram_bytes_per_launch"]
    except (OSError, KeyError, ValueError):
        return None


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def op_bytes(ops, tips, sites):
    """algorithmic HBM bytes of a streaming op list: child = 1 B/site if tip mask,
    else 128 B CLV + 4 B scaler; parent write 128 + 4 (compact tips)"""
    b = 0
    for op in ops:
        for c in (op[2], op[5]):
            b += 1 if c < tips else 132
        b += 132
    return b * sites


# --------------------------------------------------------------------------
# reference arm / CPU baseline: the unmodified reference through libpllref.so
# --------------------------------------------------------------------------
class RefWorker:
    """one private reference partition + the SingleBranchOptimizer::EvaluateMove
    call sequence for every candidate of one tree
    (src/move_tester/single_branch_optimizer.cpp:23-64)"""

    def __init__(self, R, newick, labels, seqs, radius=0, rates=None):
        self.p = R.open_mem(newick, labels, seqs, FREQS, SUBST, RATES if rates is None else rates)
        self.cands = Tree(newick).nni_candidates()
        self.radius = radius
        self.nrec = len(self.p.records())

    def run_pass_neighborhood(self):
        """BranchNeighborhoodOptimizer::EvaluateMove per candidate
        (src/move_tester/branch_neighborhood_optimizer.cpp:27-80); the clone the reference
        works on is replaced by undoing the move and restoring the changed lengths"""
        p = self.p
        t0 = time.perf_counter()
        p.traversal_update(p.root, True)
        base = [p.length(i) for i in range(self.nrec)]
        for (pidx, kind, _, _, _) in self.cands:
            p.nni(pidx, kind)
            p.invalidate_edge(pidx)
            p.traversal_update(pidx, False)
            p.optimize_branch_neighborhood(pidx, self.radius)
            p.traversal_update(pidx, False)
            p.loglikelihood(pidx)
            p.nni(pidx, kind)
            for i in range(self.nrec):
                if p.length(i) != base[i]:
                    p.update_branch_length(i, base[i])
            p.traversal_update(pidx, True)
        return time.perf_counter() - t0

    def run_pass(self, limit=None, keep=None):
        """one pass over the candidates (the first `limit`, or those with index in `keep`);
        self.ops counts the CLV operations the reference's TraversalUpdate ran
        (lib/libptpll/src/pll_partition.cpp:348-393: its return value, summed)"""
        if self.radius > 0:
            return self.run_pass_neighborhood()
        p = self.p
        self.ops = 0
        self.scores = {}
        t0 = time.perf_counter()
        self.ops += len(p.traversal_update(p.root, True))
        for i, (pidx, kind, _, _, l0) in enumerate(self.cands):
            if (limit is not None and i >= limit) or (keep is not None and i not in keep):
                continue
            p.nni(pidx, kind)
            p.invalidate_edge(pidx)
            self.ops += len(p.traversal_update(pidx, False))
            length = p.optimize_branch(pidx)
            self.scores[i] = (length, p.loglikelihood(pidx))
            p.update_branch_length(pidx, l0)
            p.nni(pidx, kind)
            p.invalidate_edge(pidx)
        return time.perf_counter() - t0

    def close(self):
        self.p.close()


def tester_name(radius):
    return ("SingleBranchOptimizer (radius 0)" if radius == 0
            else f"BranchNeighborhoodOptimizer (radius {radius})")


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_baseline_single(taxa, sites_full, aln_sample, tree, sample_sites, radius=0, rates=None, limit=None):
    from oracle.pyoracle import RefLib, have_ref
    if not have_ref():
        return None
    R = RefLib()
    labels = tree.labels()
    seqs = [bytes(aln_sample[l][:sample_sites]) for l in labels]
    w = RefWorker(R, tree.newick(), labels, seqs, radius, rates=rates)
    times = [w.run_pass(limit=limit) if radius == 0 else w.run_pass() for _ in range(3)]
    ncand = len(w.cands) if limit is None else min(limit, len(w.cands))
    w.close()
    best = min(times[1:])
    return dict(value=ncand / best * sample_sites / sites_full, unit="candidates/s", cores=1,
                kind="reference",
                sample=f"{ncand} candidates x {sample_sites} of {sites_full} sites, 1 thread, "
                       f"best of 2 passes ({best:.2f} s), scaled by sites ratio",
                raw_candidates_per_s_at_sample=ncand / best)


def parity_at_full_size(e, tree, aln, cand_rows, plan_rows, scores, n, S, k):
    """The compiled reference (oracle/_ref/libpllref.so, one thread, one ~16 GB partition at
    64 x 1M) on the bench's own tree and FULL alignment, outside every timed region: k candidates
    spread over the list scored with SingleBranchOptimizer::EvaluateMove's call sequence, and the
    per-site scaler counts of every directional CLV the batch used that the reference's FULL
    traversal also produces.  Index arithmetic beyond 2^31 bytes per buffer is exactly what a
    size-independent property cannot vouch for."""
    from oracle.pyoracle import RefLib, have_ref
    if not have_ref():
        return dict(skipped="oracle/_ref/libpllref.so missing (build() needs /root/reference)")
    labels = tree.labels()
    t0 = time.perf_counter()
    try:
        w = RefWorker(RefLib(), tree.newick(), labels, [bytes(aln[l]) for l in labels], 0)
    except MemoryError as exc:
        return dict(skipped="MemoryError: %s" % exc)
    keep = sorted(set(int(x) for x in np.linspace(0, len(cand_rows) - 1, k)))
    w.run_pass(keep=set(keep))
    lens, lnls, _ = scores
    rel_lnl = max(abs(lnls[i] - w.scores[i][1]) / abs(w.scores[i][1]) for i in keep)
    rel_len = max(abs(lens[i] - w.scores[i][0]) / max(1.0, abs(w.scores[i][0])) for i in keep)
    # scaler counts: the reference's CLV of an inner node after a FULL traversal towards the
    # root is the directional CLV of the record that faces the root
    p = w.p
    p.traversal_update(p.root, True)
    recs = p.records()
    have = set(int(r[0]) for r in plan_rows)
    facing = []

    def post(r):
        if r.is_tip:
            return
        post(r.next.back)
        post(r.next.next.back)
        facing.append(r)

    post(tree.root.back)
    post(tree.root)
    compared, mismatched, total = 0, 0, 0
    for r in facing:
        if r.node_index not in have:
            continue
        ref = p.scaler(recs[r.node_index]["scaler"])
        got = e.get_scaler(r.node_index - n)
        compared += 1
        mismatched += int(not np.array_equal(ref, got))
        total += int(ref.sum(dtype=np.int64))
    w.close()
    ok = rel_lnl <= 1e-9 and rel_len <= 1e-6 and mismatched == 0 and compared > 0
    return dict(sites=S, candidates_compared=len(keep), candidate_indices=keep, max_rel_lnl=rel_lnl,
                max_rel_len=rel_len, scaler_buffers_compared=compared, scaler_buffers_mismatched=mismatched,
                scaler_sum=total, scaler_sum_equal=mismatched == 0, passed=bool(ok),
                reference="oracle/_ref/libpllref.so (the reference compiled in place), 1 thread",
                seconds=time.perf_counter() - t0)


def cpp_partition_leg(tree, aln, n, S, lnl_sum, ms_python):
    """The same step through the C++ surface a pt build uses: pt_b200/bin/score_neighborhood builds a
    pt::pll::Partition from files and calls Partition::ScoreNniNeighborhood (what BatchedNeighborhood
    calls once per tree) on the bench's tree and FULL alignment; outside every timed region."""
    import tempfile
    exe = os.path.join(ROOT, "pt_b200", "bin", "score_neighborhood")
    info = os.path.join(ROOT, "tests", "golden", "data", "ds1", "RAxML_info.DS1")   # DS1's GTR+G4 = FREQS/SUBST/alpha
    if not os.path.exists(exe):
        return dict(skipped="pt_b200/bin/score_neighborhood not built (make -C pt_b200/host drivers REF=<pt checkout>)")
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "t.nwk"), "w") as f:
            f.write(tree.newick() + "\n")
        with open(os.path.join(d, "a.fasta"), "wb") as f:
            for label in tree.labels():
                f.write(b">" + label.encode() + b"\n" + bytes(aln[label]) + b"\n")
        r = subprocess.run([exe, os.path.join(d, "t.nwk"), os.path.join(d, "a.fasta"), info, "0", "3"],
                           capture_output=True, text=True)
    if r.returncode != 0:
        return dict(failed=r.stderr[-300:])
    out = json.loads(r.stdout)
    rel = abs(out["lnl_sum"] - lnl_sum) / abs(lnl_sum)
    out.update(call="pt::pll::Partition::ScoreNniNeighborhood (pt_b200/host), incl. P matrices + directional CLVs",
               lnl_sum_rel_diff_vs_c_abi_path=rel, ms_per_step_c_abi_path=ms_python, passed=bool(rel <= 1e-12))
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from concurrent.futures import ThreadPoolExecutor
    from oracle.pyoracle import RefLib, have_ref
    if not have_ref():
        print(json.dumps({"impl": "reference",
                          "unavailable": "oracle/_ref/libpllref.so missing (build() needs /root/reference)"}))
        return
    T = host_cores()
    sample = int(min(20000, max(2000, 16e9 / T / (2 * args.taxa * 128 * 4))))
    sample = min(sample, args.sites)
    base = parent_tree(args.taxa, 0)
    aln = simulate(base, sample, np.random.default_rng(SEED_ALN))
    R = RefLib()
    trees = [parent_tree(args.taxa, r) for r in range(T)]
    workers = [RefWorker(R, t.newick(), t.labels(), [bytes(aln[l]) for l in t.labels()], args.radius)
               for t in trees]
    ncand = sum(len(w.cands) for w in workers)
    step_s = []
    with ThreadPoolExecutor(T) as ex:
        for i in range(args.warmup + args.steps):
            t0 = time.perf_counter()
            list(ex.map(lambda w: w.run_pass(), workers))
            dt = time.perf_counter() - t0
            if i >= args.warmup:
                step_s.append(dt)
    ops_per_pass = sum(getattr(w, "ops", 0) for w in workers)
    # one thread alone on the sample (what a core does without its neighbours' memory traffic)
    alone = min(workers[0].run_pass() for _ in range(2))
    for w in workers[1:]:
        w.close()
    total = sum(step_s)
    sampled = ncand * len(step_s) / total * sample / args.sites
    # the extrapolation checked, not assumed: ONE thread at the FULL site count (the same
    # config as the GPU arm: 16 GB partition at 64 x 1M) on a bounded share of the candidates
    full = None
    if args.radius == 0 and args.sites > sample and not args.no_full_size_check:
        try:
            aln_full = simulate(base, args.sites, np.random.default_rng(SEED_ALN))
            t0 = trees[0]
            wf = RefWorker(R, t0.newick(), t0.labels(), [bytes(aln_full[l]) for l in t0.labels()], 0)
            k = min(len(wf.cands), args.full_size_candidates)
            t_full = wf.run_pass(limit=k)
            ops_full = wf.ops
            wf.close()
            # the same share on the sample, alone, for the per-site ratio
            t_small = min(workers[0].run_pass(limit=k) for _ in range(2))
            per_site_full, per_site_small = t_full / args.sites, t_small / sample
            full = dict(sites=args.sites, threads=1, candidates=k, seconds=t_full,
                        candidates_per_s=k / t_full,
                        clv_site_updates_per_s=ops_full * args.sites / t_full,
                        per_site_cost_full_over_sample=per_site_full / per_site_small,
                        note="FULL traversal + the first %d candidates of one tree; the same on the %d-site sample "
                             "took %.3f s" % (k, sample, t_small))
        except MemoryError as exc:   # a 16 GB partition did not fit this host
            full = dict(skipped="MemoryError: %s" % exc)
    workers[0].close()
    # value = the sampled all-core figure, corrected by the measured cost of a site at full size
    scale = full["per_site_cost_full_over_sample"] if full and "seconds" in full else 1.0
    value = sampled / scale
    line = dict(metric="nni_candidates_scored_per_s", value=value, unit="candidates/s",
                impl="reference", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * total / len(step_s), higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f64", data="synthetic",
                config=dict(workload=f"synthetic {args.taxa} taxa x {args.sites} sites GTR+G4, "
                                     f"all {ncand // T} NNI candidates of one tree per thread, "
                                     + tester_name(args.radius),
                            taxa=args.taxa, sites=args.sites, rate_cats=4),
                cpu_baseline=dict(value=value, unit="candidates/s", cores=T, kind="reference",
                                  sample=f"{T} threads x one tree ({ncand // T} candidates) x {sample} "
                                         f"of {args.sites} sites per step, scaled by the sites ratio and by the "
                                         f"measured per-site cost at the full size ({scale:.3f}x the sample's)"),
                # Sum of TraversalUpdate's ops_count x sites / time, all threads
                clv_site_updates_per_s=ops_per_pass * len(step_s) * sample / total / scale,
                sampled_all_cores=dict(value_scaled_by_sites_ratio_only=sampled, threads=T, sites_per_thread=sample,
                                       one_thread_alone_s_per_pass=alone,
                                       all_threads_s_per_pass=total / len(step_s)),
                full_size_one_thread=full, same_config=bool(full and "seconds" in full),
                e2e=dict(value=value, unit="candidates/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


# --------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------
def run_ours(args):
    import torch
    from pt_b200.engine import Engine, ops_array

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    sampler = ClockSampler(local)   # started long before the timed region (see stop())
    if rank == 0 and not os.environ.get("PTGPU_BENCH_NO_SAMPLER"):   # diagnostic switch only
        sampler.start()
    n, S = args.taxa, args.sites
    base = parent_tree(n, 0)
    aln = simulate(base, S, np.random.default_rng(SEED_ALN))
    tree = parent_tree(n, rank)
    labels = tree.labels()
    S_sites, weights = S, None
    if args.compress_patterns:
        # NOT the headline configuration (pt never compresses; BASELINE's shapes are in sites):
        # the same alignment as its unique columns + multiplicities, prepared once
        # outside the timed region like the alignment itself
        from pt_b200.engine import compress_site_patterns
        keys = list(aln.keys())
        cols, weights = compress_site_patterns([bytes(aln[k]) for k in keys])
        aln = {k: np.frombuffer(c.encode(), dtype=np.uint8).copy() for k, c in zip(keys, cols)}
        S = len(weights)
    # raw alignment in pinned host memory (what a caller hands to the C ABI)
    host_aln = torch.empty((n, S), dtype=torch.uint8).pin_memory()
    for i, tip in enumerate(tree.nodes[:n]):
        host_aln[tip.clv_index].copy_(torch.from_numpy(aln[tip.label]))
    nbuf = 3 * (n - 2)
    e = Engine(n, nbuf, S, 4, 2 * n - 3, nbuf, device=local)
    e.set_model(FREQS, SUBST, RATES)
    if weights is not None:
        e.set_pattern_weights(weights)

    tip_ids = np.arange(n, dtype=np.uint32)
    tip_ptrs = [host_aln[t].data_ptr() for t in range(n)]

    def upload_tips():
        e.set_tip_states_batch(tip_ids, tip_ptrs)

    upload_tips()
    e_sumtable_doubles = e.batch_sumtable_doubles()
    br = tree.branches()
    midx = np.array(list(br.keys()), dtype=np.uint32)
    mlen = np.array(list(br.values()), dtype=np.float64)
    cand_rows = tree.nni_candidates()
    cands = cand_array(cand_rows)
    ncand = len(cands)

    prog = None
    if args.radius > 0:
        from pt_b200.local_blo import Program
        prog = Program(tree, args.radius, n, nbuf, nbuf, 2 * n - 3)
    last = {}
    # only the directional CLVs the batch reads (and what they are computed from): none that
    # faces a tip, nor the one facing a cherry from its only inner neighbour
    plan_rows = tree.directional_plan(prog.referenced_clvs if prog is not None
                                      else tree.referenced_clvs(cand_rows))
    plan = ops_array(plan_rows)

    def step():
        e.update_prob_matrices(midx, mlen)
        e.update_partials(plan)
        if prog is None:
            return e.score_nni_batch(cands)
        lens, lnl, status, sm = e.optimize_local_batch(prog)
        last["smoothings"] = sm
        return lens, lnl, status

    for _ in range(args.warmup):
        lens, lnls, status = step()
    assert not status.any() and np.all(np.isfinite(lnls))

    stream = torch.cuda.ExternalStream(e.stream(), device=torch.device("cuda", local))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e.set_profiling(True)
    e.kernel_times(reset=True)
    c0 = e.counters()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    wall0 = time.time()
    ev0.record(stream)
    step_host_ms = []   # host clock per step (every step ends in a stream sync): shows WHICH step a stall hit
    for _ in range(args.steps):
        t_s = time.perf_counter()
        step()
        step_host_ms.append(round(1e3 * (time.perf_counter() - t_s), 3))
    ev1.record(stream)
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    kt = e.kernel_times(reset=True)
    c1 = e.counters()
    e.set_profiling(False)

    # end to end: raw alignment in pinned host memory -> scores on the host
    t_e2e = []
    for i in range(1 + args.steps):
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        upload_tips()
        out = step()
        torch.cuda.synchronize()
        if i:
            t_e2e.append(time.perf_counter() - t0)
    clocks = sampler.stop(wall0, time.time()) if rank == 0 else None
    e2e_s = sum(t_e2e)

    if dist:
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = float(t[0]), float(t[1])
    # parity at the headline size against the compiled reference, outside the timed regions
    parity = None
    if rank == 0 and world == 1 and args.radius == 0 and weights is None and not args.no_parity_at_size:
        parity = parity_at_full_size(e, tree, aln, cand_rows, plan_rows, out, n, S,
                                     max(3, min(args.full_size_candidates, 8)))
        if "passed" in parity and not parity["passed"]:
            raise SystemExit("parity at the full size FAILED against the reference: " + json.dumps(parity))
    # BASELINE configs[2] after the main timed region, in the same run: the same alignment with
    # its SITES sharded over the N GPUs, one tree visited (full-tree lnL + OptimizeAllBranches)
    sharded = None
    if not args.no_sharded and args.radius == 0 and weights is None:
        e.close()
        del host_aln
        sharded = sharded_visit(n, S, rank, world, local, dist, steps=min(max(args.steps, 2), 3), warmup=1,
                                aln=aln)
    cpp_leg = None
    if rank == 0 and world == 1 and args.radius == 0 and weights is None and not args.no_parity_at_size:
        if sharded is None:
            e.close()
        cpp_leg = cpp_partition_leg(tree, aln, n, S, float(np.sum(lnls)), ms / args.steps)
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return

    steps = args.steps
    value = world * ncand * steps / (ms * 1e-3)
    launches = c1["launches"] - c0["launches"]
    site_updates = c1["site_updates"] - c0["site_updates"]
    passes = c1["newton_passes"] - c0["newton_passes"]
    lnl_sweeps = c1["lnl_sweeps"] - c0["lnl_sweeps"]
    final_sweeps = c1["final_sweeps"] - c0["final_sweeps"]
    peak, peak_kind = measured_peak()
    child_bytes = sum((1 if c < n else 132) for row in cand_rows for c in (row[2][2], row[2][5], row[3][2], row[3][5]))
    st_bytes = 8 * e_sumtable_doubles     # a candidate's sumtable: 128 B/site, or 104 (compact)
    algo = dict(
        partials=op_bytes(plan_rows, n, S),                                  # per launch
        nni_prepare=(child_bytes + ncand * (st_bytes + 4)) * S,               # per launch
        # per launch: every derivative sweep reads the candidate's sumtable; the sweeps that
        # also sum the site lnL read the scaler word too; rare extra lnL-only sweeps
        newton=(passes * st_bytes + lnl_sweeps * 4 + final_sweeps * (st_bytes + 4)) / steps * S,
    )
    if prog is not None:
        # the persistent local-optimisation kernel (timed under the "newton" slot): per
        # candidate 2 prologue CLV updates + 2 edge lnL, and per executed smoothing the body's
        # CLV updates (396 B/site), sumtables (384 B/site, the first derivative evaluation rides
        # on the build), one edge lnL (268 B/site); every further derivative sweep 128 B/site.
        # Smoothings that ended on "no improvement" are not in the counter, so this is a lower
        # bound of the bytes (and of `achieved`).
        body = prog.step_types[prog.n_prologue:prog.n_prologue + prog.n_body]
        n_op, n_opt = int((body == 1).sum()), int((body == 2).sum())
        per_body = n_op * 396 + n_opt * 384 + 268
        smoothings = int(last["smoothings"].sum())
        algo = dict(partials=algo["partials"],
                    newton=(ncand * (2 * 396 + 2 * 268) + smoothings * per_body
                            + max(0.0, passes / steps - smoothings * n_opt) * 128) * S)
    kernels = {}
    for k, bytes_per_launch in algo.items():
        tot_ms, nl = kt[k]
        if nl:
            per = tot_ms / nl
            kernels[k] = dict(ms_per_launch=per, launches_per_step=nl / steps,
                              share_of_step=tot_ms / ms,
                              achieved_gbs=bytes_per_launch / (per * 1e-3) / 1e9,
                              algorithmic_bytes_per_launch=int(bytes_per_launch))
    dom = max(kernels, key=lambda k: kernels[k]["share_of_step"])
    roof = dict(bound="hbm", kernel=dom, achieved=kernels[dom]["achieved_gbs"], peak=peak,
                unit="GB/s", frac=kernels[dom]["achieved_gbs"] / peak, peak_source=peak_kind,
                traffic=ncu_traffic(dom, n, S) if prog is None else None)
    h2d = n * S + midx.nbytes + mlen.nbytes + plan.nbytes + cands.nbytes
    d2h = ncand * (8 + 8 + 4)
    line = dict(
        metric="nni_candidates_scored_per_s", value=value, unit="candidates/s",
        n_gpus=world, steps=steps, warmup=args.warmup, ms_per_step=ms / steps,
        higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64", data="synthetic",
        config=dict(workload=f"synthetic {n} taxa x {S_sites} sites GTR+G4 (BASELINE configs[2]/[3] shape)"
                             + (f", PATTERN-COMPRESSED to {S} unique columns + weights (not the headline "
                                f"configuration)" if weights is not None else "") + ": "
                             f"per step all {ncand} NNI candidates of one parent tree per GPU, "
                             + tester_name(args.radius) + " semantics, incl. P matrices and "
                             f"the {len(plan_rows)} directional CLVs the candidates read "
                             f"(of 3(n-2) = {3 * (n - 2)})", taxa=n, sites=S_sites, patterns=S, rate_cats=4,
                    candidates_per_step_per_gpu=ncand, parallelism=f"candidate-parallel x{world}",
                    l2="inputs (>20 GB of CLVs per step) far exceed the 126 MB L2; no flush needed"),
        step_host_ms=step_host_ms, e2e_step_host_ms=[round(1e3 * x, 3) for x in t_e2e],
        clv_site_updates_per_s=world * site_updates / (kt["partials"][0] * 1e-3) if kt["partials"][0] else None,
        newton_iterations_per_candidate=passes / steps / ncand,
        lnl_carrying_sweeps_per_candidate=lnl_sweeps / steps / ncand,
        extra_lnl_sweeps_per_candidate=final_sweeps / steps / ncand,
        gpu_launches=int(launches), kernels=kernels, roofline=roof, clocks=clocks,
        e2e=dict(value=world * ncand * len(t_e2e) / e2e_s, unit="candidates/s",
                 h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(d2h),
                 note="raw alignment in pinned host memory uploaded + encoded every step"),
        lnl_checksum=float(np.sum(lnls)),
    )
    if parity is not None:
        line["parity_at_size"] = parity
    if sharded is not None:
        line["sharded"] = sharded
    if cpp_leg is not None:
        line["cpp_partition"] = cpp_leg
    if world == 1 and not args.no_cpu_baseline:
        cb = cpu_baseline_single(n, S, aln, tree, min(S, args.cpu_sample_sites), args.radius)
        if cb:
            line["cpu_baseline"] = cb
    print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


# --------------------------------------------------------------------------
# BASELINE configs[3] and configs[4]: the other named shapes
# --------------------------------------------------------------------------
def run_shape(args):
    """`--workload cand100k`: BASELINE configs[3] -- 64 taxa x 100,000 sites, 10^5 NNI candidates
    per batch = the 122 candidates of each of 820 parent trees (the seeded NNI walk of SURVEY 8d),
    parents dealt round-robin to the GPUs, no collective (strong scaling: the batch is the same
    at every N).  One STEP = the whole batch: per parent P matrices, the directional CLVs its
    candidates read, and ptgpu_score_nni_batch.
    `--workload stress1000`: BASELINE configs[4] -- 1000 taxa x 50,000 sites (25.5 GB of
    directional CLVs, deep traversals, scaler counts; `--alpha 0.05` drives them up): per step
    and GPU the P matrices, a FULL traversal (998 CLV updates) + the tree's lnL, then the
    directional CLVs and all 1994 NNI candidates of the tree in one batch (weak scaling: every
    GPU its own parent tree)."""
    import torch
    from pt_b200.engine import Engine, gamma_rates_mean, ops_array

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    stress = args.workload == "stress1000"
    n = 1000 if stress else 64
    S = 50000 if stress else 100000
    parents_total = world if stress else args.parents
    rates = RATES if args.alpha is None else gamma_rates_mean(args.alpha, 4)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    base = parent_tree(n, 0)
    aln = simulate(base, S, np.random.default_rng(SEED_ALN))
    host_aln = torch.empty((n, S), dtype=torch.uint8).pin_memory()
    for tip in base.nodes[:n]:
        host_aln[tip.clv_index].copy_(torch.from_numpy(aln[tip.label]))
    nbuf = 3 * (n - 2)
    nmat = 2 * n - 3
    # P parent trees per engine call: the engine gets P sets of directional CLV / scaler buffers and
    # P sets of P matrices (plain index spaces of the C ABI: set j's indices are the tree's own
    # shifted by j x nbuf resp. j x nmat; the tips are shared), so one ptgpu_update_partials walks
    # the P trees' op lists and ONE ptgpu_score_nni_batch scores all their candidates -- the
    # Newton kernel's last iterations (a handful of slow candidates) are then shared by P x 122
    # candidates instead of 122
    P = 1 if stress else max(1, args.parents_per_call)
    e = Engine(n, P * nbuf, S, 4, P * nmat, P * nbuf, device=local)
    e.set_model(FREQS, SUBST, rates)
    tip_ids = np.arange(n, dtype=np.uint32)
    tip_ptrs = [host_aln[t].data_ptr() for t in range(n)]
    e.set_tip_states_batch(tip_ids, tip_ptrs)
    st_bytes = 8 * e.batch_sumtable_doubles()

    def shift(arr, j):
        """the op / candidate / matrix-index arrays of one tree moved into buffer set j"""
        if j == 0:
            return arr
        out = arr.copy()
        views = [out] if out.dtype.names and "parent_clv_index" in out.dtype.names else (
            [out["p"], out["q"]] if out.dtype.names else [])
        for v in views:
            for side in ("parent", "child1", "child2"):
                clv = v[side + "_clv_index"]
                clv[clv >= n] += j * nbuf
                sc = v[side + "_scaler_index"]
                sc[sc >= 0] += j * nbuf
                if side != "parent":
                    v[side + "_matrix_index"] += j * nmat
        if not views:
            out = out + np.uint32(j * nmat)
        return out

    # this rank's parents: descriptors prepared once (host integer work, like the alignment)
    work, bytes_partials, bytes_prepare, ncand_rank, full_bytes = [], 0, 0, 0, 0
    if not stress and args.parents <= 0:
        # as many parents of the walk as it takes to reach 10^5 candidates (QueueMoves' enumeration
        # yields up to 2(n-2) per tree, fewer where a node's first record faces a tip)
        parents_total, total = 0, 0
        while total < 100000:
            total += len(parent_tree(n, parents_total).nni_candidates())
            parents_total += 1
    for pidx in range(rank, parents_total, world):
        t = parent_tree(n, pidx)
        br = t.branches()
        rows = t.nni_candidates()
        plan_rows = t.directional_plan(t.referenced_clvs(rows))
        item = dict(midx=np.array(list(br.keys()), dtype=np.uint32),
                    mlen=np.array(list(br.values()), dtype=np.float64),
                    plan=ops_array(plan_rows), cands=cand_array(rows), root=t.root)
        if stress:
            ops, _, _ = t.full_traversal()
            item["full"] = ops_array(ops)
            full_bytes += op_bytes(ops, n, S) + 268 * S
        bytes_partials += op_bytes(plan_rows, n, S)
        child = sum((1 if c < n else 132) for row in rows for c in (row[2][2], row[2][5], row[3][2], row[3][5]))
        bytes_prepare += (child + len(rows) * (st_bytes + 4)) * S
        ncand_rank += len(rows)
        work.append(item)
    if P > 1:
        # groups of P parents become one item each
        grouped = []
        for g0 in range(0, len(work), P):
            grp = work[g0:g0 + P]
            grouped.append(dict(midx=np.concatenate([shift(it["midx"], j) for j, it in enumerate(grp)]),
                                mlen=np.concatenate([it["mlen"] for it in grp]),
                                plan=np.concatenate([shift(it["plan"], j) for j, it in enumerate(grp)]),
                                cands=np.concatenate([shift(it["cands"], j) for j, it in enumerate(grp)]),
                                root=grp[0]["root"]))
        work = grouped
    results = {}

    def step():
        for k, it in enumerate(work):
            e.update_prob_matrices(it["midx"], it["mlen"])
            if stress:
                e.update_partials(it["full"])
                r, rb = it["root"], it["root"].back
                results["lnl"] = e.edge_loglikelihood(r.clv_index, r.scaler_index, rb.clv_index,
                                                      rb.scaler_index, r.pmatrix_index)
            e.update_partials(it["plan"])
            results[k] = e.score_nni_batch(it["cands"])

    for _ in range(args.warmup):
        step()
    for k in range(len(work)):
        assert not results[k][2].any() and np.all(np.isfinite(results[k][1]) | np.isneginf(results[k][1]))
    stream = torch.cuda.ExternalStream(e.stream(), device=torch.device("cuda", local))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e.set_profiling(True)
    e.kernel_times(reset=True)
    c0 = e.counters()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    wall0 = time.time()
    ev0.record(stream)
    for _ in range(args.steps):
        step()
    ev1.record(stream)
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    kt = e.kernel_times(reset=True)
    c1 = e.counters()
    e.set_profiling(False)
    t_e2e = []
    for i in range(1 + min(args.steps, 2)):
        if dist:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e.set_tip_states_batch(tip_ids, tip_ptrs)
        step()
        torch.cuda.synchronize()
        if i:
            t_e2e.append(time.perf_counter() - t0)
    clocks = sampler.stop(wall0, time.time()) if rank == 0 else None
    e2e_s = sum(t_e2e) / len(t_e2e)
    tot = torch.tensor([float(ncand_rank), float(c1["site_updates"] - c0["site_updates"])],
                       dtype=torch.float64, device="cuda")
    if dist:
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = float(t[0]), float(t[1])
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ncand_all, site_updates = float(tot[0]), float(tot[1])
    if rank != 0:
        if dist:
            dist.destroy_process_group()
        return
    steps = args.steps
    passes = c1["newton_passes"] - c0["newton_passes"]
    lnl_sweeps = c1["lnl_sweeps"] - c0["lnl_sweeps"]
    final_sweeps = c1["final_sweeps"] - c0["final_sweeps"]
    peak, peak_kind = measured_peak()
    # rank 0's kernels, totals over the timed steps; algorithmic bytes as in the headline workload
    algo = dict(partials=(bytes_partials + (full_bytes - 268 * S * len(work) if stress else 0)) * steps,
                nni_prepare=bytes_prepare * steps,
                newton=(passes * st_bytes + lnl_sweeps * 4 + final_sweeps * (st_bytes + 4)) * S,
                edge_lnl=268 * S * len(work) * steps if stress else 0)
    kernels = {}
    for k, b in algo.items():
        k_ms, k_n = kt[k]
        if k_n:
            kernels[k] = dict(ms_total_per_step=k_ms / steps, launches_per_step=k_n / steps,
                              share_of_step=k_ms / ms, achieved_gbs=b / (k_ms * 1e-3) / 1e9,
                              algorithmic_bytes_per_step=int(b / steps))
    dom = max(kernels, key=lambda k: kernels[k]["share_of_step"])
    value = ncand_all * steps / (ms * 1e-3)
    h2d = n * S + sum(it["midx"].nbytes + it["mlen"].nbytes + it["plan"].nbytes + it["cands"].nbytes
                      + (it["full"].nbytes if stress else 0) for it in work)
    name = ("synthetic 1000 taxa x 50000 sites GTR+G4 (BASELINE configs[4]): per step and GPU P matrices, FULL "
            "traversal (998 CLV updates) + lnL, the directional CLVs and all 1994 NNI candidates of one tree"
            if stress else
            f"synthetic 64 taxa x 100000 sites GTR+G4 (BASELINE configs[3]): per step a batch of {int(ncand_all)} NNI "
            f"candidates = all 122 of each of {parents_total} parent trees (seeded NNI walk), parents dealt "
            f"round-robin to the GPUs, {P} parents' candidates per engine call")
    line = dict(metric="nni_candidates_scored_per_s", value=value, unit="candidates/s", n_gpus=world, steps=steps,
                warmup=args.warmup, ms_per_step=ms / steps, higher_is_better=True,
                scaling="weak" if stress else "strong", vs_baseline=None, dtype="f64", data="synthetic",
                config=dict(workload=name, taxa=n, sites=S, rate_cats=4, alpha=args.alpha,
                            candidates_per_step=int(ncand_all), parents_per_step=parents_total, parents_per_call=P,
                            parallelism=f"candidate-parallel x{world}",
                            l2="every parent's CLVs and sumtables (> 1 GB) exceed the 126 MB L2; no flush needed"),
                clv_site_updates_per_s=site_updates / (ms * 1e-3),
                newton_iterations_per_candidate=passes / steps / max(1, ncand_rank),
                gpu_launches=int(c1["launches"] - c0["launches"]), kernels=kernels,
                roofline=dict(bound="hbm", kernel=dom, achieved=kernels[dom]["achieved_gbs"], peak=peak,
                              unit="GB/s", frac=kernels[dom]["achieved_gbs"] / peak, peak_source=peak_kind,
                              traffic=None),
                clocks=clocks,
                e2e=dict(value=ncand_all / e2e_s, unit="candidates/s", h2d_bytes_per_step=int(h2d),
                         d2h_bytes_per_step=int(ncand_rank * 20 + (8 if stress else 0)),
                         note="raw alignment uploaded + encoded once per step, descriptors per parent; bytes per rank"),
                lnl_checksum=float(sum(np.sum(results[k][1][np.isfinite(results[k][1])]) for k in range(len(work)))))
    if stress:
        line["tree_lnl"] = results["lnl"]
    if world == 1 and not args.no_cpu_baseline:
        t0 = parent_tree(n, 0)
        sample = 1000 if stress else 20000
        cb = cpu_baseline_single(n, S, aln, t0, sample, 0, rates=rates)
        if cb:
            line["cpu_baseline"] = cb
    print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


# --------------------------------------------------------------------------
# BASELINE configs[1] on DS1: the real drivers, wall clock
# --------------------------------------------------------------------------
def run_ptw_ds1(args):
    ds1 = os.path.join(ROOT, "tests", "golden", "data", "ds1")
    import tempfile

    def walk(binary, threads):
        with tempfile.TemporaryDirectory() as d:
            out = os.path.join(d, "good.tsv")
            tested = os.path.join(d, "tested.tsv")
            cmd = [binary, "--raxml-info", os.path.join(ds1, "RAxML_info.DS1"), "--threads", str(threads),
                   "--polling-time", "1", "--lnl-offset", "-5", "--optimization-radius", "1", "--tested-trees", tested,
                   os.path.join(ds1, "RAxML_bestTree.DS1"), os.path.join(ds1, "DS1.fasta"), out]
            t0 = time.perf_counter()
            env = dict(os.environ)
            if binary == ours:
                env["PTGPU_HOST_STATS"] = "1"   # the driver's own wall-clock marks of its phases (stderr)
            if binary == ours and args.gpus > 0:
                # wanderer-parallel over the first N GPUs: every wanderer's Partition takes the next
                # visible device round-robin (pt_b200/host/src/pll_partition.cpp, pick_device)
                env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(args.gpus))
            r = subprocess.run(cmd, capture_output=True, text=True, check=True, env=env)
            dt = time.perf_counter() - t0
            good = sum(1 for _ in open(out))
            ntested = sum(1 for _ in open(tested))
            moves = None
            phases = {}
            for line in r.stderr.splitlines():
                if line.startswith("batched move testing:"):
                    moves = int(line.split()[3])
                if line.startswith("[ptw_batched]"):
                    phases[line[14:-10].strip()] = float(line.split()[-2])
            return dt, good, ntested, moves, phases

    ours = os.path.join(ROOT, "pt_b200", "bin", "ptw_batched")
    ref = os.path.join(ROOT, "oracle", "_ref", "ptw_threads")
    if not os.path.exists(ours):
        raise SystemExit("pt_b200/bin/ptw_batched is not built (make -C pt_b200/host drivers REF=<pt checkout>)")
    sampler = ClockSampler(0)
    sampler.start()
    W = max(1, args.wanderers)
    for _ in range(args.warmup):
        walk(ours, W)
    wall0 = time.time()
    runs = [walk(ours, W) for _ in range(args.steps)]
    clocks = sampler.stop(wall0, time.time())
    dt = sum(r[0] for r in runs) / len(runs)
    _, good, ntested, moves, _ = runs[-1]
    # where the process's wall time went, mean over the timed runs (CUDA start-up -- cuInit + the
    # primary context, 0.5-2.4 s from box to box -- is inside "guru"; the rest of the process's
    # wall time is the dynamic loader and process exit)
    phases = {k: round(sum(r[4].get(k, 0.0) for r in runs) / len(runs), 3) for k in runs[-1][4]}
    line = dict(metric="nni_candidates_scored_per_s", value=moves / dt, unit="candidates/s", n_gpus=args.gpus,
                steps=args.steps, warmup=args.warmup, ms_per_step=1e3 * dt, higher_is_better=True,
                scaling="strong", vs_baseline=None, dtype="f64", data="DS1 (27 taxa x 1949 sites)",
                config=dict(workload="BASELINE configs[1] on DS1: ptw topography walk, --lnl-offset -5 "
                                     "--optimization-radius 1, %d wanderer(s) with one engine each, dealt round-robin "
                                     "to %d GPU(s); batched move testing, visits ahead of time (whole process, incl. "
                                     "CUDA start-up: see process_phases_s)" % (W, args.gpus),
                            wanderers=W,
                            taxa=27, sites=1949, rate_cats=4),
                good_trees=good, tested_trees=ntested, evaluate_move_calls=moves, clocks=clocks,
                process_phases_s=phases,
                e2e=dict(value=moves / dt, unit="candidates/s", h2d_bytes_per_step=27 * 1949,
                         d2h_bytes_per_step=8 * moves,
                         note="the run IS end to end: files in, tables out"))
    if os.path.exists(ref) and not args.no_cpu_baseline:
        T = host_cores()
        rdt, rgood, rtested, _, _ = walk(ref, T)
        line["cpu_baseline"] = dict(value=moves / rdt, unit="candidates/s", cores=T, kind="reference",
                                    sample=f"the unmodified ptw_threads --threads {T}, same arguments, "
                                           f"{rdt:.1f} s wall, {rgood} good / {rtested} tested trees "
                                           "(thread interleaving makes the multi-wanderer counts vary by a few)")
    print(json.dumps(line))


# --------------------------------------------------------------------------
# BASELINE configs[2]: one tree, sites sharded over the GPUs (strong scaling)
# --------------------------------------------------------------------------
def sharded_reference_leg(taxa, sites, sample, steps=1):
    """the unmodified reference on one host thread (a single tree gives libpll nothing to run
    in parallel): FULL traversal + lnL + OptimizeAllBranches on `sample` sites, scaled"""
    from oracle.pyoracle import RefLib, have_ref
    if not have_ref():
        return None
    tree = parent_tree(taxa, 0)
    aln = simulate(tree, sample, np.random.default_rng(SEED_ALN))
    labels = tree.labels()
    best = None
    for _ in range(1 + steps):
        p = RefLib().open_mem(tree.newick(), labels, [bytes(aln[l]) for l in labels], FREQS, SUBST, RATES)
        t0 = time.perf_counter()
        p.traversal_update(p.root, True)
        p.loglikelihood(p.root)
        p.optimize_all_branches(p.root)
        p.traversal_update(p.root, False)
        lnl = p.loglikelihood(p.root)
        dt = time.perf_counter() - t0
        p.close()
        best = dt if best is None else min(best, dt)
    return dict(value=1.0 / (best * sites / sample), unit="trees/s", cores=1, kind="reference",
                sample=f"FULL traversal + lnL + OptimizeAllBranches of one {taxa}-taxon tree on {sample} of "
                       f"{sites} sites, 1 thread (one tree is sequential in the reference), best of "
                       f"{1 + steps} ({best:.2f} s), scaled by sites ratio", seconds_at_sample=best)


def sharded_visit(n, S, rank, world, local, dist, steps, warmup, call_by_call=False, exchange="peer",
                  parity=True, aln=None):
    """BASELINE configs[2]: one tree, the SITES split over the ranks (contiguous tile-aligned
    ranges).  One unit = what Wanderer::MoveForward does for a tree it visits
    (src/wanderer.cpp:150-210): FULL TraversalUpdate, LogLikelihood, OptimizeAllBranches,
    LogLikelihood -- four engine calls per rank, the whole-tree optimisation being ONE persistent
    kernel per GPU that sums d1/d2 and lnL over the shards itself (NVLink peer memory, rank
    order).  Strong scaling: same tree and alignment at every N.  Returns rank 0's record (None
    elsewhere): ms per tree (CUDA events on the engine's stream, max over ranks), the same
    procedure unsharded on rank 0's GPU in the same run (=> efficiency), launches / exchanges per
    tree, and the parity verdict against the unsharded run (scaler counts integer-identical,
    lnL within 1e-12)."""
    import hashlib
    import torch
    from pt_b200 import all_branches as ab
    from pt_b200.engine import Engine
    from pt_b200.sharded import SiteShardedEngine, shard_range

    tree = parent_tree(n, 0)
    if aln is None:
        aln = simulate(tree, S, np.random.default_rng(SEED_ALN))
    newick = tree.newick()
    tip_ids = np.arange(n, dtype=np.uint32)

    def pinned_rows(b0, b1):
        host = torch.empty((n, b1 - b0), dtype=torch.uint8).pin_memory()
        for tip in tree.nodes[:n]:
            host[tip.clv_index].copy_(torch.from_numpy(np.ascontiguousarray(aln[tip.label][b0:b1])))
        return host, [host[t].data_ptr() for t in range(n)]

    def scaler_digests(e, ranges):
        """sha256 over all n-2 scaler rows, one digest per site range"""
        hs = [hashlib.sha256() for _ in ranges]
        for i in range(n - 2):
            row = e.get_scaler(i)
            for h, (b, e_) in zip(hs, ranges):
                h.update(np.ascontiguousarray(row[b:e_]).tobytes())
        return [h.hexdigest() for h in hs]

    def timed(e, stream_dev, barrier):
        info = {}

        def unit():
            t = Tree(newick)              # every unit starts from the input lengths
            info["lnl0"], info["lnl"], info["passes"] = ab.visit(e, t, call_by_call=call_by_call)
            info["lengths"] = t.branches()

        for _ in range(warmup):
            unit()
        stream = torch.cuda.ExternalStream(e.stream(), device=torch.device("cuda", stream_dev))
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0 = e.counters()
        barrier()
        torch.cuda.synchronize()
        ev0.record(stream)
        for _ in range(steps):
            unit()
        ev1.record(stream)
        torch.cuda.synchronize()
        barrier()
        c1 = e.counters()
        info["ms"] = ev0.elapsed_time(ev1) / steps
        info["launches"] = (c1["launches"] - c0["launches"]) / steps
        info["site_updates"] = (c1["site_updates"] - c0["site_updates"]) / steps
        info["newton_passes"] = (c1["newton_passes"] - c0["newton_passes"]) / steps
        return info

    bar = (lambda: dist.barrier()) if (dist and world > 1) else (lambda: None)
    b0, b1 = shard_range(S, world, rank)
    if world > 1:
        e = SiteShardedEngine(n, n - 2, S, 4, 2 * n - 3, n - 2, device=local, exchange=exchange)
        eng = e.engine
    else:
        e = eng = Engine(n, n - 2, S, 4, 2 * n - 3, n - 2, device=local)
    e.set_model(FREQS, SUBST, RATES)
    host, ptrs = pinned_rows(b0, b1)
    eng.set_tip_states_batch(tip_ids, ptrs)
    x0 = e.exchanges() if world > 1 else 0
    info = timed(e, local, bar)
    exchanges = ((e.exchanges() - x0) / (warmup + steps)) if world > 1 else 0
    digest = scaler_digests(eng, [(0, b1 - b0)])[0] if parity else None
    ms = info["ms"]
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
        upd = torch.tensor([info["site_updates"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(upd, op=dist.ReduceOp.SUM)
        site_updates = float(upd[0])
        digests = [None] * world
        dist.all_gather_object(digests, digest)
    else:
        site_updates, digests = info["site_updates"], [digest]
    e.close()
    del host
    rec = None
    if rank == 0:
        rec = dict(ms_per_tree=ms, trees_per_s=1e3 / ms, n_gpus=world, sites_per_gpu=b1 - b0,
                   procedure=("call by call" if call_by_call else
                              "one persistent step-program kernel per GPU (ptgpu_optimize_all_branches)"),
                   exchange=(("in-kernel sums over NVLink peer memory, rank order" if exchange == "peer"
                              else "torch.distributed allreduce hook") if world > 1 else "none (one GPU)"),
                   gpu_launches_per_tree=info["launches"], engine_calls_per_tree=4 if not call_by_call else None,
                   host_round_trips_inside_optimize_all_branches=0 if not call_by_call else None,
                   cross_gpu_exchanges_per_tree=exchanges,
                   nvlink_payload_bytes_per_tree_per_gpu=int(exchanges * 2 * 16 * max(0, world - 1)),
                   optimize_all_branches_passes=info["passes"], newton_evaluations_per_tree=info["newton_passes"],
                   clv_site_updates_per_s=site_updates / (ms * 1e-3),
                   lnl_start=info["lnl0"], lnl_optimised=info["lnl"])
        if world > 1 and parity:
            # the same procedure unsharded on this rank's GPU, in the same run
            e1 = Engine(n, n - 2, S, 4, 2 * n - 3, n - 2, device=local)
            e1.set_model(FREQS, SUBST, RATES)
            host, ptrs = pinned_rows(0, S)
            e1.set_tip_states_batch(tip_ids, ptrs)
            one = timed(e1, local, lambda: None)
            ranges = [shard_range(S, world, r) for r in range(world)]
            same = scaler_digests(e1, ranges) == digests
            e1.close()
            rel = abs(info["lnl"] - one["lnl"]) / abs(one["lnl"])
            rel0 = abs(info["lnl0"] - one["lnl0"]) / abs(one["lnl0"])
            len_rel = max(abs(info["lengths"][k] - v) / max(1e-8, abs(v)) for k, v in one["lengths"].items())
            rec.update(ms_per_tree_1gpu_same_run=one["ms"],
                       strong_scaling_efficiency=one["ms"] / (world * ms),
                       speedup_vs_1gpu=one["ms"] / ms,
                       parity=dict(scaler_counts_identical=bool(same), lnl_rel_diff=rel, lnl_start_rel_diff=rel0,
                                   passes_equal=info["passes"] == one["passes"], max_length_rel_diff=len_rel,
                                   passed=bool(same and rel <= 1e-12 and rel0 <= 1e-12)))
    if world > 1:
        dist.barrier()
    return rec


def run_sharded(args):
    """`--workload sharded`: BASELINE configs[2] as the line's own metric (trees visited/s, strong
    scaling); the default workload carries the same record as its `sharded` sub-record."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            T = host_cores()
            cb = sharded_reference_leg(args.taxa, args.sites, min(args.sites, 20000), args.steps)
            if cb is None:
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libpllref.so missing"}))
                return
            print(json.dumps(dict(metric="trees_visited_per_s", value=cb["value"], unit="trees/s", impl="reference",
                                  n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                                  ms_per_step=1e3 / cb["value"], higher_is_better=True, scaling="strong",
                                  vs_baseline=None, dtype="f64", data="synthetic",
                                  config=dict(workload=f"synthetic {args.taxa} taxa x {args.sites} sites GTR+G4: "
                                              "full-tree lnL + OptimizeAllBranches of one tree",
                                              taxa=args.taxa, sites=args.sites, rate_cats=4, host_cores=T),
                                  cpu_baseline=cb,
                                  e2e=dict(value=cb["value"], unit="trees/s", h2d_bytes_per_step=0,
                                           d2h_bytes_per_step=0))))
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    wall0 = time.time()
    rec = sharded_visit(args.taxa, args.sites, rank, world, local, dist, args.steps, args.warmup,
                        call_by_call=args.call_by_call, exchange=args.exchange)
    clocks = sampler.stop(wall0, time.time()) if rank == 0 else None
    if rank == 0:
        n, S = args.taxa, args.sites
        line = dict(metric="trees_visited_per_s", value=rec["trees_per_s"], unit="trees/s", n_gpus=world,
                    steps=args.steps, warmup=args.warmup, ms_per_step=rec["ms_per_tree"], higher_is_better=True,
                    scaling="strong", vs_baseline=None, dtype="f64", data="synthetic",
                    config=dict(workload=f"synthetic {n} taxa x {S} sites GTR+G4 (BASELINE configs[2]): per step FULL "
                                         "traversal + lnL + OptimizeAllBranches + lnL of ONE tree, sites sharded over "
                                         "the GPUs (contiguous tile-aligned ranges), per-tree scalars summed between "
                                         "the GPUs only", taxa=n, sites=S, rate_cats=4,
                                parallelism=f"site-sharded x{world}",
                                l2="every pass streams this rank's CLVs (>= 16 MB each): no flush needed"),
                    gpu_launches=int(rec["gpu_launches_per_tree"] * args.steps), sharded=rec, clocks=clocks,
                    e2e=dict(value=rec["trees_per_s"], unit="trees/s", h2d_bytes_per_step=0, d2h_bytes_per_step=16,
                             note="alignment resident; the visit's scalars come back to the host"))
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = sharded_reference_leg(n, S, min(S, 20000))
        print(json.dumps(line))
    if dist:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--taxa", type=int, default=64)
    ap.add_argument("--sites", type=int, default=1000000)
    ap.add_argument("--radius", type=int, default=0,
                    help="0: SingleBranchOptimizer semantics (the headline workload); >= 1: "
                         "BranchNeighborhoodOptimizer semantics (BASELINE configs[1]'s tester)")
    ap.add_argument("--cpu-sample-sites", type=int, default=100000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-full-size-check", action="store_true",
                    help="reference arm: skip the one-thread run at the full site count")
    ap.add_argument("--full-size-candidates", type=int, default=24,
                    help="reference arm / parity at size: candidates the reference scores at the full site count")
    ap.add_argument("--no-parity-at-size", action="store_true",
                    help="skip the comparison with the compiled reference at the full site count")
    ap.add_argument("--workload", default="synthetic",
                    choices=["synthetic", "ptw-ds1", "sharded", "cand100k", "stress1000"])
    ap.add_argument("--parents", type=int, default=0,
                    help="cand100k: parent trees per batch (0 = as many as give 10^5 candidates)")
    ap.add_argument("--parents-per-call", type=int, default=8,
                    help="cand100k: parent trees whose candidates are scored by ONE engine call")
    ap.add_argument("--alpha", type=float, default=None,
                    help="cand100k / stress1000: Gamma shape instead of DS1's 0.134946 (0.05: scaler stress)")
    ap.add_argument("--compress-patterns", action="store_true",
                    help="score the alignment's unique columns with weights (SURVEY 8f-4; not the headline)")
    ap.add_argument("--call-by-call", action="store_true",
                    help="sharded: OptimizeAllBranches as the reference's call sequence instead of one kernel")
    ap.add_argument("--exchange", default="peer", choices=["peer", "hook"],
                    help="sharded: in-kernel peer-memory sums (default) or the NCCL allreduce hook")
    ap.add_argument("--no-sharded", action="store_true",
                    help="default workload: skip the site-sharded sub-record (BASELINE configs[2])")
    ap.add_argument("--wanderers", type=int, default=1,
                    help="ptw-ds1: wanderer threads of ptw_batched (each owns an engine on the GPU)")
    args = ap.parse_args()
    if args.workload == "ptw-ds1":
        if int(os.environ.get("RANK", "0")) == 0:
            run_ptw_ds1(args)
    elif args.workload == "sharded":
        run_sharded(args)
    elif args.workload in ("cand100k", "stress1000"):
        if args.impl == "reference":
            raise SystemExit("--impl reference covers the default workload and --workload sharded; the other "
                             "shapes carry the reference as their cpu_baseline leg")
        run_shape(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
