#!/usr/bin/env python
"""Benchmark of the self-play hot path (BASELINE.json: MCTS simulations/s and self-play games/s).

A "step" is one move search for every concurrent game -- n_trees x `iters` simulations of the reference's loop
(ai/mcts.py:16) incl. the policy/value net on every leaf that is not answered by the evaluation cache -- followed by
the device-side move (csrc/az_selfplay.cu).  Successive steps are successive plies of the SAME games (slots of
finished games are refilled), which is what self-play is and what lets the evaluation cache (MCTS.cache) see the
cross-move reuse the reference gets.  Default workload = BASELINE.json configs[1]: Connect4 6x7, 4096 concurrent
games, 800 sims/move, virtual-loss batch 1 (the reference's sequential search), random-init main.py:134 net (D=512,
patch 5, 8 heads, depth 1), games starting from synthetic positions of 0..7 random plies.  configs[2] (Gomoku 7x7,
8192 games, vl = 8) is measured in the same run and reported under `workloads.gomoku`.

  value     device-timed (CUDA events on the launching stream), everything resident in HBM
  e2e       wall clock through the public API with HOST buffers, every step: pinned float32 root planes + noise in,
            visit counts + root Q out, the move applied by the caller on the host (what a user of MCTS.mcts_batch does)
  selfplay  whole games from the empty board to the end on every rank + the NCCL all-gather of the trajectories inside
            the timed region: games/s at N GPUs
  --impl reference   the reference's CPU implementation of the same path on the host cores (oracle/cpu_baseline.py):
            the oracle port for Connect4 (the reference cannot run it with a network), the UNMODIFIED reference from
            baseline/_ref for Gomoku; 1-process and P-process figures.

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    "connect4": dict(game="connect4", rows=6, trees=4096, iters=800, vl=1, max_plies=7,
                     name="connect4_6x7_4096games_800sims_vl1_net512x1"),
    "gomoku": dict(game="gomoku", rows=7, trees=8192, iters=800, vl=8, max_plies=12,
                   name="gomoku_7x7_8192games_800sims_vl8_net512x1"),
    # the reference's other network (main.py:187-188: D = 256, depth 2): the generic evaluator path (all-token tcgen05 attention,
    # full-block GEMMs); not part of the default run, `--workload gomoku_depth2`
    "gomoku_depth2": dict(game="gomoku", rows=7, trees=4096, iters=800, vl=8, max_plies=12, net=dict(dim=256, depth=2, patch=5, heads=8),
                          name="gomoku_7x7_4096games_800sims_vl8_net256x2"),
}
NET = dict(dim=512, depth=1, patch=5, heads=8)          # main.py:134
ROUNDS_CONTINUOUS = 4                                   # vl = 1, continuous plies: simulations per tree and launch while leaves hit the cache


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=250, help="timed steps (plies); the default gives a >= 20 s timed region")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="connect4", choices=sorted(WORKLOADS))
    ap.add_argument("--trees", type=int, default=0)
    ap.add_argument("--iters", type=int, default=0)
    ap.add_argument("--vl", type=int, default=0)
    ap.add_argument("--min-seconds", type=float, default=0.0, help="keep stepping until the timed region is at least this long")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the e2e leg (default min(steps, 48))")
    ap.add_argument("--cpu-seconds", type=float, default=0.0, help="per CPU figure (default 15 in the b200 arm, 30 in the reference arm)")
    ap.add_argument("--cpu-procs", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-selfplay", action="store_true", help="skip the whole-game self-play leg (games/s)")
    ap.add_argument("--no-second-workload", action="store_true", help="skip the workloads.<other> sub-record")
    ap.add_argument("--no-profile", action="store_true")
    ap.add_argument("--no-cache", action="store_true", help="evaluation cache off (A/B)")
    ap.add_argument("--lockstep", action="store_true", help="vl = 1: search all slots together ply by ply (A/B against continuous plies)")
    ap.add_argument("--rounds", type=int, default=0, help="vl = 1: cached simulations per tree and launch in continuous self-play (0 = default)")
    ap.add_argument("--no-train", action="store_true", help="skip the training leg (config 5)")
    ap.add_argument("--train-iters", type=int, default=40, help="timed optimiser steps of the training leg")
    ap.add_argument("--selfplay-generations", type=float, default=6.0, help="whole games per slot in the self-play leg")
    return ap.parse_args()


# ---- CPU baseline (subprocess workers; the only place bench.py runs oracle/) -------------------------------
def have_ref():
    return os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "games"))


def run_cpu_baseline(wl, seconds, procs, kind="auto"):
    """P single-thread worker processes side by side.  kind: 'reference' (unmodified reference from baseline/_ref,
    Gomoku 7x7 only), 'port' (oracle port), 'auto' = reference where it can run."""
    ncpu = os.cpu_count() or 1
    procs = procs or min(ncpu, 32)
    if kind == "auto":
        kind = "reference" if (wl["game"] == "gomoku" and wl["rows"] == 7 and have_ref()) else "port"
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "cpu_baseline.py"), "--kind", kind, "--game", wl["game"], "--rows", str(wl["rows"]),
           "--dim", str(wl.get("net", NET)["dim"]), "--depth", str(wl.get("net", NET)["depth"]), "--patch", str(wl.get("net", NET)["patch"]),
           "--heads", str(wl.get("net", NET)["heads"]),
           "--iters", str(wl["iters"]), "--seconds", str(seconds)]
    env = dict(os.environ, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    t0 = time.time()
    ps = [subprocess.Popen(cmd + ["--seed", str(i)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env, text=True)
          for i in range(procs)]
    rate, sims, hits, ok, err = 0.0, 0, [], 0, ""
    for p in ps:
        so, se = p.communicate()
        out = so.strip().splitlines()
        if p.returncode == 0 and out:
            r = json.loads(out[-1])
            rate += r["sims"] / r["seconds"]
            sims += r["sims"]
            hits.append(r.get("cache_hit_rate", 0.0))
            ok += 1
        else:
            err = (se or "").strip().splitlines()[-1:] or [""]
            err = err[0][:200]
    what = ("the UNMODIFIED reference (baseline/_ref: Gomoku().self_play with its own Net, MCTS.cache on, counted by MCTS.mcts_count)"
            if kind == "reference" else
            "the oracle port of ai/mcts.py (per-node Python objects, batch-1 torch Net on CPU, MCTS.cache analogue)")
    d = dict(value=rate, unit="simulations/s", cores=ok, kind=kind,
             sample=f"{ok} single-thread processes x {seconds:.0f}s of whole self-play games, {what}, {wl['game']} {wl['iters']} sims/move, "
                    f"Dirichlet on; {sims} sims total; host has {ncpu} logical cores",
             cache_hit_rate=(sum(hits) / len(hits)) if hits else None, wall_s=time.time() - t0)
    if ok < procs:
        d["worker_errors"] = f"{procs - ok} of {procs} workers failed: {err}"
    return d


# ---- clocks ---------------------------------------------------------------------------------
class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                       "-i", str(gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def summary(self, t0, t1):
        if self.p:
            self.p.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 8] or [r for _, r in self.rows if len(r) >= 8]
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        sm = sorted(int(float(r[1])) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[4 + i].lower().startswith("active") for r in rows)]
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=int(float(rows[0][2])), reasons=reasons, samples=len(rows),
                    power_w_max=max(float(r[3]) for r in rows))


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(tflops=p["bf16_tflops_sustained"], hbm=p["hbm_gbs"], src_tensor="measured (MEASURED_PEAKS.json, sustained bf16: kernels timed inside a long step)",
                    src_hbm="measured (MEASURED_PEAKS.json, HBM copy bandwidth)")
    except Exception:
        return dict(tflops=1400.0, hbm=6650.0, src_tensor="fallback (B200_PROFILING.md: ~1.4 PFLOP/s sustained bf16)",
                    src_hbm="fallback (B200_PROFILING.md: 6.65 TB/s)")


def ncu_traffic(kind, wl):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` capture of this
    kernel on this workload (profiles/ncu_traffic.json names the capture file), else None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)
        return t[kind][wl["name"]]["bytes"]
    except Exception:
        return None


def roofline(wl, r, prof, pk, kind=None):
    """Roofline entry of one kernel kind (default: the kind with the largest SHARE of the step; the three cls-row GEMM
    launches form one kind).  Algorithmic work per launch = per-leaf (or per-simulation) figure of DESIGN.md section 4 x
    the units of one launch; duration = CUDA-event average over the profiled launches."""
    G = game_class(wl)
    net = wl.get("net", NET)
    D, H, P = net["dim"], net["heads"], net["patch"]
    T = G.rows * G.cols + 1
    kused = G.feature_dim * P * P + T
    leaves = r["leaves_per_step"]
    adv_n = max(prof["forest_advance"]["launches"], 1)
    per_step = {k: v["ms"] * (v["launches"] / adv_n if k != "forest_advance" else 1.0) for k, v in prof.items()}
    if kind is None:
        kind = max(per_step, key=lambda k: per_step[k])
    ms = prof[kind]["ms"]
    traffic = ncu_traffic(kind, wl)
    st = r["stats"]
    share = per_step[kind] / sum(per_step.values())
    flops, ref_flops = None, None
    front2 = bool(r.get("path_flags", 0) & 2)
    depth = net["depth"]
    lps = max(prof[kind]["launches"] / adv_n, 1e-9)       # launches of this kind per search pass (the generic path evaluates in chunks of 4096 boards)
    if kind == "front_fused" and front2:   # statistics (H scores + K Cholesky columns) per token + U = A^T w: H columns per token
        flops, name = 2.0 * leaves * T * kused * (H + kused + H), "az_front2_kernel (embedding+LN1+score projection+cls softmax+U=A^T w, tcgen05)"
    elif kind == "front_fused":     # one contraction over the K binary inputs per token: D value + H score + K Gram columns
        flops, name = 2.0 * leaves * T * kused * (D + H + kused), "az_front_kernel (embedding+LN1+V/score projection+cls attention, tcgen05)"
    elif kind == "gemm_kv":
        flops, name = 2.0 * leaves * T * D * r["proj_width"] / lps, f"az_gemm_kernel (all-token projection, N={r['proj_width']}, tcgen05)"
    elif kind == "gemm_cls_rows":   # out_proj D*D (D x H*128 after the front-kernel reassociation), mlp D*4D, 4D*D over 3 launches -> average launch
        kout = H * 128 if front2 else D
        qproj = 0 if r.get("path_flags", 0) & 3 else D * D        # generic path: the cls query projection is a fourth launch
        flops, name = 2.0 * leaves * (kout * D + 8 * D * D + qproj) / lps, "az_gemm_kernel (cls-row out_proj / MLP GEMMs, tcgen05; average over the launches of a pass)"
        ref_flops = 2.0 * leaves * (D * D + 8 * D * D + qproj) / lps          # the reference's graph: out-proj over K = D
    elif kind == "tail_fused":      # AZ_FUSED_TAIL=1: the three cls-row GEMMs + the head GEMM in one cluster kernel
        kout = H * 128 if front2 else D
        flops, name = 2.0 * leaves * (kout * D + 8 * D * D + 64 * D), "az_tail_kernel (out-proj + MLP + heads of the cls rows, one cluster kernel, tcgen05)"
    elif kind == "embed_ln":
        flops, name = 2.0 * leaves * T * 128 * D / lps, "az_embed_mma_kernel (im2col embedding + LN, tcgen05)"
    elif kind == "gemm_full_block":  # per full block: in-proj D x 3D, out-proj D x D, mlp D x 4D and 4D x D over all tokens
        flops, name = 2.0 * leaves * T * 12 * D * D * (depth - 1) / lps, "az_gemm_kernel (full-block GEMMs over all tokens, tcgen05 + TMA stores; average over the launches of a pass)"
    elif kind == "attention_full":   # S = Q K^T and O = P V per head: 2 x 2 T^2 hd per head
        flops, name = 4.0 * leaves * T * T * D * (depth - 1) / lps, "az_attn_mma_kernel (all-token attention of a full block, tcgen05)"
    if flops is not None:
        ach = flops / (ms / 1e3) / 1e12
        d = dict(kernel=name, bound="tensor", achieved=ach, peak=pk["tflops"], unit="TFLOP/s", frac=ach / pk["tflops"],
                 traffic=traffic, peak_source=pk["src_tensor"], flops_per_launch=flops, ms_per_launch=ms, share_of_step=share,
                 rows_per_launch=leaves / max(lps / {"gemm_cls_rows": 3.0 if r.get("path_flags", 0) & 3 else 4.0, "gemm_full_block": 4.0 * max(depth - 1, 1),
                                                     "attention_full": float(max(depth - 1, 1))}.get(kind, 1.0), 1.0))
        if ref_flops is not None:
            d["reference_graph"] = dict(flops_per_launch=ref_flops, achieved=ref_flops / (ms / 1e3) / 1e12,
                                        frac=ref_flops / (ms / 1e3) / 1e12 / pk["tflops"],
                                        note="out-proj counted over K = D as in ai/nn.py; the executed graph contracts over K = heads*128")
        return d
    if kind == "forest_advance":    # 16 B per child scored + 8 B header per select + 24 B RMW per path node + 26 B per child written
        sims = max(st["sims"], 1)
        per_sim = (16.0 * st["child_reads"] + 8.0 * st["selects"] + 24.0 * (st["selects"] + st["sims"]) + 26.0 * st["children"]) / sims
        sims_per_launch = r.get("sims_per_advance", wl["trees"] * wl["vl"])
        nbytes = per_sim * sims_per_launch
        ach = nbytes / (ms / 1e3) / 1e9
        return dict(kernel="az_advance_kernel (PUCT select + expand + backup + cache probe, one warp per tree)", bound="hbm", achieved=ach,
                    peak=pk["hbm"], unit="GB/s", frac=ach / pk["hbm"], traffic=traffic, peak_source=pk["src_hbm"], bytes_per_sim=per_sim,
                    sims_per_launch=sims_per_launch, bytes_per_launch=nbytes, ms_per_launch=ms, share_of_step=share)
    return None


def game_class(wl):
    from alpha_zero_b200 import games
    G = {"connect4": games.Connect4, "gomoku": games.Gomoku, "tictactoe": games.TicTacToe}[wl["game"]]
    if wl["game"] == "gomoku":
        G.rows = G.cols = wl["rows"]
        G.action_dim = G.state_dim = wl["rows"] ** 2
    return G


def host_ply(G, az_train, np, boards, players, plies, visits):
    """What a caller of MCTS.mcts_batch does between two searches (e2e leg): the most visited action per game, make_move,
    check_winner; finished games restart from the empty board.  Vectorised NumPy (train._moves_from_actions / _wins)."""
    acts = np.argmax(visits, axis=1)
    r, c = az_train._moves_from_actions(G, boards, acts)
    idx = np.arange(len(acts))
    movers = players.astype(np.int64)
    boards[idx, movers, r, c] = 1
    if G.feature_dim == 3:
        boards[:, 2] = (1 - movers)[:, None, None]
    players[:] = 1 - movers
    plies += 1
    over = az_train._wins(G, boards, movers, r, c) | (plies == G.state_dim)
    boards[over] = 0
    players[over] = 0
    plies[over] = 0
    return int(over.sum())


def run_workload(args, wl, rank, world, steps, warmup, full):
    """All legs of one workload on this rank.  full: also e2e bookkeeping details, profile pass (rank 0), clocks."""
    import numpy as np
    import torch
    import torch.distributed as dist
    import alpha_zero_b200 as az
    from alpha_zero_b200 import parallel, train as az_train
    from alpha_zero_b200.ai.nn import Net, B200Net
    from alpha_zero_b200.selfplay import SelfPlayEngine, reference_sample_plies
    from alpha_zero_b200.synth import random_roots

    local = int(os.environ.get("LOCAL_RANK", 0))
    G = game_class(wl)
    n, iters, vl = wl["trees"], wl["iters"], wl["vl"]
    A = G.action_dim
    torch.manual_seed(0)                                              # identical random-init weights on every rank
    net = wl.get("net", NET)
    model = Net((G.rows, G.cols), net["patch"], net["dim"], A, net["heads"], net["depth"], G.feature_dim, dropout=0.1).eval()
    stream = torch.cuda.Stream()
    eng = B200Net(model, max_batch=n * vl)
    forest = az.Forest(az.game_spec(wl["game"], wl["rows"]), n, iters, vl=vl, stream=stream.cuda_stream)
    if args.no_cache:
        forest.cache_config(0)
    rng = np.random.default_rng(1234 + rank)
    sample_plies = reference_sample_plies(G, True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(*vals):
        if world == 1:
            return list(vals)
        tm = torch.tensor(vals, device="cuda", dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        return tm.tolist()

    out = dict(workload=wl["name"], cache=forest.cache_info())
    clocks = ClockSampler(local) if full else None
    # ---- value: device time, everything resident; continuous device-side self-play -------------------------
    # K "steps" = K x iters search passes (one pass = advance + evaluator over all concurrent games); a slot whose
    # search is finished moves and starts its next search at once, so the simulations done are counted, not assumed.
    total = warmup + steps
    planes, players, mcs = random_roots(G, n, wl["max_plies"], rng)
    forest.set_roots_planes(planes, players, mcs)                     # upload BEFORE the timed region
    continuous = vl == 1 and not args.lockstep                        # measured: continuous plies gain 46 % at vl = 1 and lose 7 % at vl = 8
    rounds = args.rounds or (ROUNDS_CONTINUOUS if continuous else 1)
    forest.set_rounds(rounds)
    n_games = int(n * (4 + total))                                    # enough games to refill every finished slot
    sp = SelfPlayEngine(forest, n_games, 1234, sample_plies, noise=True, id_first=rank, id_stride=world)
    sp.begin(adopt_roots=True)
    ev, passes = [], 0

    def lockstep_plies(k):                                            # all slots search together, then all move
        for _ in range(k):
            sp.prepare()                                              # Dirichlet noise of the new roots (device)
            forest.search_net(eng, iters)                             # n x iters simulations
            sp.move()                                                 # record, choose, apply, refill (device)
    if continuous:
        sp.run(eng, iters, max_steps=warmup * iters)                  # warm-up: graph capture, caches, clocks
    else:
        lockstep_plies(warmup)
    barrier()
    forest.stats(reset=True)
    t_clock0 = time.time()
    launches0 = eng.launch_count() + sp.launch_count()
    while True:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        if continuous:
            passes += sp.run(eng, iters, max_steps=steps * iters, resume=True)
        else:
            lockstep_plies(steps)
        b.record(stream)
        ev.append((a, b))
        torch.cuda.synchronize()
        if args.min_seconds <= 0 or sum(x.elapsed_time(y) for x, y in ev) >= args.min_seconds * 1e3 or len(ev) >= 64:
            break
    barrier()
    steps_timed = steps * len(ev)
    if not continuous:
        passes = None
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    gpu_launches = eng.launch_count() + sp.launch_count() - launches0
    stats = forest.stats(reset=True)
    st = sp.status()
    assert st["live"] == n, f"self-play ran out of games ({st})"
    assert stats["sims"] >= 0.5 * n * iters * steps_timed, (stats["sims"], n * iters * steps_timed)
    sp.close()
    dev_ms, = max_over_ranks(dev_ms)
    sims_all = float(stats["sims"])
    if world > 1:
        tsum = torch.tensor([sims_all], device="cuda", dtype=torch.float64)
        dist.all_reduce(tsum)
        sims_all = float(tsum.item())
    out.update(dev_ms=dev_ms, steps=steps_timed, sims_timed=sims_all, search_passes=passes, rounds=rounds, continuous=continuous,
               gpu_launches=int(gpu_launches),
               stats=stats, games_finished_in_value_leg=st["finished"])
    forest.set_rounds(1)                                              # the per-ply API below searches in lockstep
    # ---- e2e: wall clock through the public API with host buffers --------------------------------------------
    e2e_steps = args.e2e_steps or min(steps, 48)
    pinned = torch.empty((n, G.feature_dim, G.rows, G.cols), dtype=torch.float32).pin_memory()
    boards = pinned.numpy()                                           # the caller's boards live in pinned host memory
    np.copyto(boards, planes)
    pl, mc = players.astype(np.int32).copy(), mcs.astype(np.int32).copy()
    noise = [rng.dirichlet([0.03] * A, size=n) for _ in range(4)]
    forest.set_active(None)
    parts = []

    def e2e_step(k):
        ta = time.perf_counter()
        forest.set_roots_planes(boards, pl, mc)                       # H2D inside the timed region
        forest.set_noise(noise[k % 4])                                # H2D
        tb = time.perf_counter()
        forest.search_net(eng, iters)
        tc = time.perf_counter()
        visits, _q = forest.root_visits()                             # D2H of the step's result
        td = time.perf_counter()
        host_ply(G, az_train, np, boards, pl, mc, visits)
        parts.append((tb - ta, tc - tb, td - tc, time.perf_counter() - td))
        return visits
    for k in range(min(warmup, 3)):
        e2e_step(k)
    barrier()
    parts.clear()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        visits = e2e_step(k)
    barrier()
    e2e_s = time.perf_counter() - t0
    assert int(visits.sum()) == n * (iters - 1), "every tree must have completed its simulations"
    e2e_s, = max_over_ranks(e2e_s)
    t_clock1 = time.time()
    med = lambda xs: sorted(xs)[len(xs) // 2] * 1e3
    out.update(e2e_s=e2e_s, e2e_steps=e2e_steps, e2e_sims_per_step=n * iters * world, h2d=int(pinned.numel() * 4 + n * 8 + n * A * 8), d2h=int(n * A * 4 + n * 8),
               e2e_step_ms=dict(upload=round(med([p[0] for p in parts]), 3), search=round(med([p[1] for p in parts]), 3),
                                readback=round(med([p[2] for p in parts]), 3), host_move=round(med([p[3] for p in parts]), 3)),
               e2e_stats=forest.stats(reset=True))
    if clocks:
        out["clocks"] = clocks.summary(t_clock0, t_clock1)
    # ---- whole games on every rank + the NCCL all-gather of the trajectories inside the timed region ---------
    if not args.no_selfplay:
        per_rank = max(int(n * args.selfplay_generations), 1)
        games_total = per_rank * world
        sp2 = SelfPlayEngine(forest, per_rank, 4321, sample_plies, noise=True, id_first=rank, id_stride=world)
        forest.set_rounds(rounds)
        forest.stats(reset=True)
        # host side of the call: the model's weights live in host memory (a torch state_dict), the trajectories end in pinned host arrays
        host_sd = {k: v.detach().cpu() for k, v in model.state_dict().items()}
        g_spec = forest.g
        P_, A_ = g_spec.state_dim, g_spec.action_dim
        pinned_out = dict(plies=torch.empty(games_total, dtype=torch.int32).pin_memory(), winner=torch.empty(games_total, dtype=torch.int32).pin_memory(),
                          first=torch.empty(games_total, dtype=torch.int32).pin_memory(),
                          words=torch.empty((games_total, P_, 2, g_spec.nw), dtype=torch.int64).pin_memory(),
                          visits=torch.empty((games_total, P_, A_), dtype=torch.int16).pin_memory(),
                          q=torch.empty((games_total, P_), dtype=torch.float64).pin_memory(),
                          action=torch.empty((games_total, P_), dtype=torch.int16).pin_memory())
        barrier()
        t0 = time.perf_counter()
        eng.load_state_dict(host_sd)                                  # H2D: weights -> tensor-core engine (new weights: cold evaluation cache)
        tw = time.perf_counter()
        sp2.begin()
        searches = sp2.run(eng, iters, lockstep=not continuous)       # device-side self-play until every game is over
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        merged = parallel.gather_trajectories(sp2.device_tensors(), games_total)     # one all_gather_into_tensor (NCCL)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        for k_, dst in pinned_out.items():                            # D2H: every rank reads all games (what fills its replay buffer)
            dst.copy_(merged[k_], non_blocking=True)
        torch.cuda.synchronize()
        t2b = time.perf_counter()
        barrier()
        t3 = time.perf_counter()
        positions = int(merged["plies"].sum().item()) if world > 1 else sp2.status()["positions"]
        assert int((merged["winner"] == -2).sum().item()) == 0, "every gathered game must be finished"
        st2 = forest.stats(reset=True)
        play_s, gather_s, total_s, upload_s, d2h_s = max_over_ranks(t1 - tw, t2 - t1, t3 - t0, tw - t0, t2b - t2)
        h2d_bytes = int(sum(v.numel() * v.element_size() for v in host_sd.values()))
        d2h_bytes = int(sum(v.numel() * v.element_size() for v in pinned_out.values()))
        forest.set_rounds(1)
        out["selfplay"] = dict(games=games_total, games_per_rank=per_rank, concurrent_per_gpu=n, positions=positions, search_passes_per_rank=searches,
                               cached_sims_per_launch=rounds,
                               seconds=total_s, play_seconds=play_s, gather_seconds=gather_s, gather_share=gather_s / total_s,
                               weight_upload_seconds=upload_s, d2h_seconds=d2h_s, h2d_bytes=h2d_bytes, d2h_bytes=d2h_bytes,
                               gather_bytes_per_rank=int(sum(v.numel() * v.element_size() for k, v in merged.items()) // world),
                               games_per_s=games_total / total_s, positions_per_s=positions / total_s,
                               simulations_per_s=positions * iters / total_s, sims_per_move=iters,
                               cache_hit_rate=st2["cache_hits"] / max(st2["expansions"], 1),
                               evaluator_rows_per_sim=(st2["expansions"] - st2["cache_hits"]) / max(st2["sims"], 1),
                               how="selfplay.SelfPlayEngine: every game from the empty board to its end on the device (Dirichlet noise, "
                                   "moves sampled / first-max as the reference; continuous plies: a slot moves when ITS search is finished; "
                                   "slots refilled until all games have started, so the batch thins out only at the end), then "
                                   "parallel.gather_trajectories: one all_gather_into_tensor of the compact "
                                   "trajectories; max over ranks")
        # ---- config 5: the gathered games -> device replay buffer -> train.train with the gradient all-reduce ---------
        if not args.no_train:
            import copy
            from alpha_zero_b200.replay_buffer import DeviceReplayBuffer
            dbuf = DeviceReplayBuffer(G, 4_000_000, seed=100 + rank)      # every rank draws its own batches
            torch.cuda.synchronize()
            ta = time.perf_counter()
            added = dbuf.add_trajectories(merged)                     # z, canonical boards, dihedral variants: one kernel
            torch.cuda.synchronize()
            add_s = time.perf_counter() - ta
            tmodel = copy.deepcopy(model)
            B, T = 512, args.train_iters                              # main.py:153 batch, train.py:93 lr below
            az_train.train(tmodel, B, dbuf, 3, 2.5e-4, "cuda")        # warm-up (cuDNN / cuBLAS plans, NCCL channels)
            barrier()
            tb = time.perf_counter()
            losses = az_train.train(tmodel, B, dbuf, T, 2.5e-4, "cuda")
            torch.cuda.synchronize()
            barrier()
            train_s, = max_over_ranks(time.perf_counter() - tb)
            ts = dict(az_train.last_stats)
            tu = time.perf_counter()
            eng.load_state_dict(tmodel.state_dict())                  # new weights into the inference engine (clears the evaluation cache)
            torch.cuda.synchronize()
            upload_s = time.perf_counter() - tu
            out["train"] = dict(samples_per_s=B * world * T / train_s, iterations=T, batch_per_gpu=B, seconds=train_s,
                                ms_per_iteration=train_s / T * 1e3, buffer_entries=dbuf.size(), entries_added=added,
                                entry_bytes=dbuf.entry_bytes(), add_entries_per_s=added / max(add_s, 1e-9),
                                grad_bytes=ts.get("grad_bytes"), allreduce_buckets=ts.get("buckets", 0),
                                allreduce_exposed_share=(ts.get("allreduce_exposed_ms", 0.0) / 1e3) / train_s,
                                weight_upload_s=upload_s, loss=losses[0], policy_loss=losses[1], value_loss=losses[2],
                                how="replay_buffer.DeviceReplayBuffer.add_trajectories on the all-gathered games, then train.train "
                                    "(train.py:85-123: Adam lr 2.5e-4, soft-target CE + MSE fused kernel + 1e-4 L2, batch 512 per GPU "
                                    "sampled on the device, bucketed NCCL all-reduce overlapped with backward), then the weights are "
                                    "uploaded to the tensor-core engine; torch (cuDNN / cuBLAS) runs the Net's forward and backward")
            dbuf.close()
            del tmodel
        del merged, pinned_out
        sp2.close()
    # ---- per-kernel profile (rank 0): eager steps of one search with CUDA events ------------------------------
    if rank == 0 and full and not args.no_profile:
        planes, players, mcs = random_roots(G, n, wl["max_plies"], rng)
        forest.set_active(None)
        forest.set_roots_planes(planes, players, mcs)
        forest.set_noise(None)
        forest.stats(reset=True)
        forest.cache_clear()                                          # the training leg changed the weights: step-by-step callers say so
        forest.set_rounds(rounds)                                     # the launches profiled are the launches timed above
        forest.begin_move(iters)
        adv, lv, per_step = [], [], {}
        eng.profile(True)
        eng.profile(False)                                            # zero the totals, timing off
        prev = {}
        for s in range(2 * iters):                                    # the whole search, every 8th step timed
            timed = s % 8 == 4
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            eng.profile_on(timed)
            a.record(stream)
            forest.advance()
            b.record(stream)
            eng.eval_leaves(forest, stream.cuda_stream)
            if timed:
                adv.append((a, b))
                nl, nd = forest.counts()
                lv.append(nl)
                cur = eng.profile_read()                              # per-step deltas: steps beyond 3x the median are dropped,
                for k, (ms, cnt) in cur.items():                      # so one preempted launch cannot skew a kernel's figure
                    pms, pcnt = prev.get(k, (0.0, 0))
                    if cnt > pcnt:
                        per_step.setdefault(k, []).append(((ms - pms) / (cnt - pcnt), cnt - pcnt))
                prev = cur
                if nd >= n:
                    break
        eng.profile(False)
        n_adv = s + 1

        def robust_mean(xs):                                          # mean over the steps within 3x the median
            m = sorted(xs)[len(xs) // 2]
            keep = [x for x in xs if x <= 3.0 * m]
            return sum(keep) / len(keep)
        pst = forest.stats(reset=True)
        out["profile"] = {k: dict(ms=robust_mean([m for m, _ in v]), launches=sum(c for _, c in v)) for k, v in per_step.items()}
        out["profile"]["forest_advance"] = dict(ms=robust_mean([a.elapsed_time(b) for a, b in adv]), launches=len(adv))
        out["leaves_per_step"] = sum(lv) / max(len(lv), 1)
        out["sims_per_advance"] = pst["sims"] / max(n_adv, 1)
        out["profile_stats"] = pst
        out["proj_width"] = eng.proj_width()
        out["path_flags"] = eng.path_flags()
    forest.close()
    eng.close()
    return out


def summarize(args, wl, r, world, pk):
    """JSON fields of one workload from run_workload's raw numbers."""
    value = r["sims_timed"] / (r["dev_ms"] / 1e3)
    e2e = r["e2e_sims_per_step"] * r["e2e_steps"] / r["e2e_s"]
    st, es = r["stats"], r["e2e_stats"]
    per_ply = dict(value=e2e, unit="simulations/s", h2d_bytes_per_step=r["h2d"], d2h_bytes_per_step=r["d2h"], steps=r["e2e_steps"],
                   step_ms_median=r["e2e_step_ms"], cache_hit_rate=es["cache_hits"] / max(es["expansions"], 1),
                   how="MCTS.mcts_batch-style use: every ply the caller uploads its float32 root planes + Dirichlet noise from pinned host "
                       "memory, searches all games in lockstep, reads visit counts + root Q back and applies the moves on the host")
    if "selfplay" in r:                                               # the call a user of the reference makes: collect_data(model) -> games
        sp_ = r["selfplay"]
        e2e_rec = dict(value=sp_["simulations_per_s"], unit="simulations/s", h2d_bytes_per_step=sp_["h2d_bytes"], d2h_bytes_per_step=sp_["d2h_bytes"],
                       steps=1, games_per_s=sp_["games_per_s"],
                       how="one call = train.collect_data's work through the public API: the model's state_dict is uploaded from host memory, "
                           f"{sp_['games']} whole games are played on the device(s), the trajectories are all-gathered and copied to pinned host "
                           "memory; wall clock over all of it, max over ranks (details under `selfplay`)", per_ply_api=per_ply)
    else:
        e2e_rec = per_ply
    d = dict(workload=wl["name"], value=value, unit="simulations/s", ms_per_step=r["dev_ms"] / r["steps"], steps=r["steps"],
             e2e=e2e_rec,
             gpu_launches=r["gpu_launches"], search_stats=st, search_passes=r["search_passes"], cached_sims_per_launch=r["rounds"],
             plies="continuous" if r["continuous"] else "lockstep",
             sims_per_pass_per_game=(st["sims"] / r["search_passes"] / wl["trees"]) if r["search_passes"] else None,
             eval_cache=dict(entries=r["cache"]["entries"], bytes=r["cache"]["bytes"], hit_rate=st["cache_hits"] / max(st["expansions"], 1),
                             evaluator_rows_per_sim=(st["expansions"] - st["cache_hits"]) / max(st["sims"], 1),
                             note="MCTS.cache analogue (ai/mcts.py:37-51): hits = leaves answered without an evaluator row"))
    if "clocks" in r:
        d["clocks"] = r["clocks"]
    if "selfplay" in r:
        d["selfplay"] = r["selfplay"]
    if "train" in r:
        d["train"] = r["train"]
    prof = r.get("profile", {})
    if prof:
        rr = dict(r, stats=r["profile_stats"])
        d["kernel_ms"] = {k: round(v["ms"], 4) for k, v in prof.items()}
        d["roofline"] = roofline(wl, rr, prof, pk)
        d["rooflines_all"] = {k: rl for k, rl in ((k, roofline(wl, rr, prof, pk, k)) for k in prof) if rl}
        pf = r.get("path_flags", 0)
        d["evaluator_path"] = dict(flags=pf, front_kernel="az_front2_kernel" if pf & 2 else ("az_front_kernel" if pf & 1 else "generic"),
                                   cls_row_gemm="cta_pair" if pf & 4 else "single_cta", layernorm_fold=bool(pf & 8), head_epilogue=bool(pf & 16),
                                   gemm_stores="tma_store" if pf & 128 else "per_thread", fused_tail=bool(pf & 64))
    return d


def config_of(wl, world):
    return dict(workload=wl["name"], game=wl["game"], board=f"{wl['rows']}x{7 if wl['game'] == 'connect4' else wl['rows']}",
                concurrent_games_per_gpu=wl["trees"], sims_per_move=wl["iters"], vl_batch=wl["vl"],
                net=("Net(D=512,patch=5,heads=8,depth=1) random-init, ai/nn.py / main.py:134" if "net" not in wl else
                     f"Net(D={wl['net']['dim']},patch={wl['net']['patch']},heads={wl['net']['heads']},depth={wl['net']['depth']}) random-init, ai/nn.py / main.py:187"),
                steps="device-side self-play from synthetic positions of 0..max_plies random plies, finished games restart from the empty board, "
                      "Dirichlet noise on, moves sampled / first-max as the reference.  vl 1: continuous plies, one step = sims_per_move search "
                      "passes (advance + evaluator over all concurrent games), a slot moves when its own search is finished; vl > 1: lockstep, "
                      "one step = one ply of every game.  value = simulations completed (device counters) / device time",
                l2="inputs larger than L2: forest node pools + activations >> 126 MB, every step searches new root positions",
                parallelism=f"dp{world} (games sharded per GPU; no collective inside a search; selfplay leg: one NCCL all-gather of trajectories)")


def main():
    args = parse()
    wl = dict(WORKLOADS[args.workload])
    for k in ("trees", "iters", "vl"):
        if getattr(args, k):
            wl[k] = getattr(args, k)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    other = None if args.no_second_workload else dict(WORKLOADS["gomoku" if wl["game"] == "connect4" else "connect4"])
    base = dict(metric="mcts_simulations_per_second", unit="simulations/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic", config=config_of(wl, world))

    if args.impl == "reference":
        if rank != 0:
            return
        secs = args.cpu_seconds or 30.0
        procs = args.cpu_procs or min(os.cpu_count() or 1, 64)
        many = run_cpu_baseline(wl, secs, procs)
        one = run_cpu_baseline(wl, secs, 1)
        line = dict(base, impl="reference", value=many["value"], ms_per_step=wl["trees"] * wl["iters"] * world / max(many["value"], 1e-9) * 1e3,
                    cpu_baseline=many, cpu_baseline_1proc=one,
                    e2e=dict(value=many["value"], unit="simulations/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0), gpu_launches=0,
                    note="ms_per_step is extrapolated (one step = concurrent games x sims/move simulations); the CPU arm times a bounded sample")
        if other:
            line["workloads"] = {other["game"]: dict(workload=other["name"], cpu_baseline=run_cpu_baseline(other, secs, procs),
                                                     cpu_baseline_1proc=run_cpu_baseline(other, secs, 1))}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = run_cpu_baseline(wl, args.cpu_seconds or 15.0, args.cpu_procs)
    r = run_workload(args, wl, rank, world, args.steps, args.warmup, full=True)
    r2 = None
    if other:
        from alpha_zero_b200.ai import MCTS
        MCTS.release()
        r2 = run_workload(args, other, rank, world, min(args.steps, 40), min(args.warmup, 3), full=True)
    if world > 1:
        dist.destroy_process_group()
    if rank != 0:
        return
    pk = peaks()
    line = dict(base, impl="b200")
    line.update(summarize(args, wl, r, world, pk))
    line["steps"] = r["steps"]
    if r2:
        line["workloads"] = {other["game"]: dict(summarize(args, other, r2, world, pk), config=config_of(other, world))}
    if cb:
        line["cpu_baseline"] = cb
    print(json.dumps(line))


if __name__ == "__main__":
    main()
