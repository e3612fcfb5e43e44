#!/usr/bin/env python
"""Benchmark of the self-play hot path (BASELINE.json: MCTS simulations/s).

A "step" is one move search for every concurrent game: n_trees x `iters` simulations of the
reference's loop (ai/mcts.py:16) incl. the policy/value net on every leaf.  Default workload
= BASELINE.json configs[1]: Connect4 6x7, 4096 concurrent games, 800 sims/move, virtual-loss
batch 1 (the reference's sequential search), random-init main.py:134 net (D=512, patch 5,
8 heads, depth 1).  `--workload gomoku` runs configs[2] (Gomoku 7x7, 8192 games, vl=8).

  value  device-timed (CUDA events on the launching stream), root positions already in HBM
  e2e    wall clock through the C ABI with HOST buffers: pinned float32 root planes in,
         visit counts + root Q out, every step
  --impl reference   the reference's CPU implementation of the same path (the oracle port of
         its single-threaded Python search + batch-1 torch net), one process per host core.

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
}
NET = dict(dim=512, depth=1, patch=5, heads=8)          # main.py:134


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="connect4", choices=sorted(WORKLOADS))
    ap.add_argument("--trees", type=int, default=0)
    ap.add_argument("--iters", type=int, default=0)
    ap.add_argument("--vl", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--cpu-procs", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-selfplay", action="store_true", help="skip the whole-game self-play leg (games/s)")
    return ap.parse_args()


# ---- CPU baseline (oracle port, subprocess workers; the only place bench.py runs oracle/) ----
def run_cpu_baseline(wl, seconds, procs):
    ncpu = os.cpu_count() or 1
    procs = procs or min(ncpu, 32)
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "cpu_baseline.py"), "--game", wl["game"], "--rows", str(wl["rows"]),
           "--dim", str(NET["dim"]), "--depth", str(NET["depth"]), "--patch", str(NET["patch"]), "--heads", str(NET["heads"]),
           "--iters", str(wl["iters"]), "--seconds", str(seconds)]
    env = dict(os.environ, OMP_NUM_THREADS="1", MKL_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    t0 = time.time()
    ps = [subprocess.Popen(cmd + ["--seed", str(i)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, env=env, text=True)
          for i in range(procs)]
    rate, sims = 0.0, 0
    for p in ps:
        out = p.communicate()[0].strip().splitlines()
        if p.returncode == 0 and out:
            r = json.loads(out[-1])
            rate += r["sims"] / r["seconds"]
            sims += r["sims"]
    return dict(value=rate, unit="simulations/s", cores=procs, kind="port",
                sample=f"{procs} single-thread processes x {seconds:.0f}s of the same search (oracle port of ai/mcts.py "
                       f"+ batch-1 torch Net on CPU, {wl['iters']} sims/move, Dirichlet on); {sims} sims total; "
                       f"host has {ncpu} logical cores", wall_s=time.time() - t0)


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
        return dict(tflops=p["bf16_tflops_sustained"], hbm=p["hbm_gbs"], src="measured (MEASURED_PEAKS.json, sustained bf16)")
    except Exception:
        return dict(tflops=1400.0, hbm=6650.0, src="fallback (B200_PROFILING.md: ~1.4 PF sustained, 6.65 TB/s)")


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
    """Roofline entry of one kernel kind (default: the kind with the largest share of the step).
    Algorithmic work per launch = per-leaf (or per-simulation) figure of DESIGN.md section 4 x the units of
    one launch; duration = CUDA-event average over the profiled launches."""
    G = game_class(wl)
    D, H, P = NET["dim"], NET["heads"], NET["patch"]
    T = G.rows * G.cols + 1
    kused = G.feature_dim * P * P + T
    leaves = r["leaves_per_step"]
    per_step = {k: v["ms"] * (v["launches"] / max(prof["forest_advance"]["launches"], 1) if k != "forest_advance" else 1.0)
                for k, v in prof.items()}
    if kind is None:                # dominant KERNEL: longest single launch (the three cls-row GEMMs are separate kernels)
        kind = max(prof, key=lambda k: prof[k]["ms"])
    ms = prof[kind]["ms"]
    traffic = ncu_traffic(kind, wl)
    st = r["stats"]
    flops = None
    front2 = bool(r.get("path_flags", 0) & 2)
    if kind == "front_fused" and front2:   # statistics (H scores + K Cholesky columns) per token + U = A^T w: H columns per token
        flops, name = 2.0 * leaves * T * kused * (H + kused + H), "az_front2_kernel (embedding+LN1+score projection+cls softmax+U=A^T w, tcgen05)"
    elif kind == "front_fused":     # one contraction over the K binary inputs per token: D value + H score + K Gram columns
        flops, name = 2.0 * leaves * T * kused * (D + H + kused), "az_front_kernel (embedding+LN1+V/score projection+cls attention, tcgen05)"
    elif kind == "gemm_kv":
        flops, name = 2.0 * leaves * T * D * r["proj_width"], f"az_gemm_kernel (all-token projection, N={r['proj_width']}, tcgen05)"
    elif kind == "gemm_cls_rows":   # out_proj D*D (D x H*128 after the front-kernel reassociation), mlp D*4D, 4D*D over 3 launches -> average launch
        kout = H * 128 if front2 else D
        flops, name = 2.0 * leaves * (kout * D + 8 * D * D) / 3.0, "az_gemm_kernel (cls-row out_proj / MLP GEMMs, tcgen05; average of the 3 launches)"
    elif kind == "embed_ln":
        flops, name = 2.0 * leaves * T * 128 * D, "az_embed_mma_kernel (im2col embedding + LN, tcgen05)"
    if flops is not None:
        ach = flops / (ms / 1e3) / 1e12
        return dict(kernel=name, bound="tensor", achieved=ach, peak=pk["tflops"], unit="TFLOP/s", frac=ach / pk["tflops"],
                    traffic=traffic, peak_source=pk["src"], flops_per_launch=flops, ms_per_launch=ms, share_of_step=per_step[kind] / sum(per_step.values()))
    if kind == "forest_advance":    # 16 B per child scored + 8 B header per select + 24 B RMW per path node + 26 B per child written
        sims = max(st["sims"], 1)
        per_sim = (16.0 * st["child_reads"] + 8.0 * st["selects"] + 24.0 * (st["selects"] + st["sims"]) + 26.0 * st["children"]) / sims
        nbytes = per_sim * wl["trees"] * wl["vl"]
        ach = nbytes / (ms / 1e3) / 1e9
        return dict(kernel="az_advance_kernel (PUCT select + expand + backup, one warp per tree)", bound="hbm", achieved=ach, peak=pk["hbm"],
                    unit="GB/s", frac=ach / pk["hbm"], traffic=traffic, peak_source=pk["src"], bytes_per_sim=per_sim, bytes_per_launch=nbytes,
                    ms_per_launch=ms, share_of_step=per_step[kind] / sum(per_step.values()))
    return None


def game_class(wl):
    from alpha_zero_b200 import games
    G = {"connect4": games.Connect4, "gomoku": games.Gomoku, "tictactoe": games.TicTacToe}[wl["game"]]
    if wl["game"] == "gomoku":
        G.rows = G.cols = wl["rows"]
        G.action_dim = G.state_dim = wl["rows"] ** 2
    return G


def run_b200(args, wl, rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist
    import alpha_zero_b200 as az
    from alpha_zero_b200.ai.nn import Net, B200Net
    from alpha_zero_b200.synth import random_roots

    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    G = game_class(wl)
    n, iters, vl = wl["trees"], wl["iters"], wl["vl"]
    A, T = G.action_dim, G.rows * G.cols + 1
    torch.manual_seed(0)                                              # identical random-init weights on every rank
    model = Net((G.rows, G.cols), NET["patch"], NET["dim"], A, NET["heads"], NET["depth"], G.feature_dim, dropout=0.1).eval()
    stream = torch.cuda.Stream()
    eng = B200Net(model, max_batch=n * vl)
    forest = az.Forest(az.game_spec(wl["game"], wl["rows"]), n, iters, vl=vl, stream=stream.cuda_stream)
    rng = np.random.default_rng(1234 + rank)                          # Philox-free: host NumPy, per-rank stream
    total = args.warmup + args.steps
    roots = [random_roots(G, n, wl["max_plies"], rng) for _ in range(total)]
    noise = [rng.dirichlet([0.03] * A, size=n) for _ in range(total)]
    pinned = torch.empty((n, G.feature_dim, G.rows, G.cols), dtype=torch.float32).pin_memory()
    pin_np = pinned.numpy()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    e2e_parts = []

    def one_step(i, e2e):
        planes, players, mcs = roots[i]
        ta = time.perf_counter()
        if e2e:
            np.copyto(pin_np, planes)                                 # the step's inputs start in pinned host memory
            forest.set_roots_planes(pin_np, players, mcs)             # H2D inside the timed region
        forest.set_noise(noise[i])
        tb = time.perf_counter()
        forest.search_net(eng, iters)
        tc = time.perf_counter()
        if e2e:
            out = forest.root_visits()                                # D2H of the step's result
            e2e_parts.append((round((tb - ta) * 1e3, 2), round((tc - tb) * 1e3, 2), round((time.perf_counter() - tc) * 1e3, 2)))
            return out

    clocks = ClockSampler(local)
    # ---- value: device time, inputs resident --------------------------------------------------
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = 0
    for i in range(total):
        planes, players, mcs = roots[i]
        forest.set_roots_planes(planes, players, mcs)                 # upload BEFORE the timed region
        forest.set_noise(noise[i])
        if i == args.warmup:
            barrier()
            t_clock0 = time.time()
            launches0 = eng.launch_count()
        torch.cuda.synchronize()
        if i >= args.warmup:
            ev[i - args.warmup][0].record(stream)
        forest.search_net(eng, iters)
        if i >= args.warmup:
            ev[i - args.warmup][1].record(stream)
    barrier()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    gpu_launches = eng.launch_count() - launches0 + args.steps        # + begin_move kernels
    stats = forest.stats(reset=True)
    # ---- e2e: wall clock through the public API with host buffers ------------------------------
    for i in range(args.warmup):
        one_step(i, True)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.warmup, total):
        visits, q = one_step(i, True)
    barrier()
    e2e_s = time.perf_counter() - t0
    t_clock1 = time.time()
    assert int(visits.sum()) == n * (iters - 1), "every tree must have completed its simulations"
    if world > 1:
        tm = torch.tensor([dev_ms, e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)                      # max over ranks
        dev_ms, e2e_s = tm.tolist()
    sims_per_step = n * iters * world
    out = dict(dev_ms=dev_ms, e2e_s=e2e_s, sims_per_step=sims_per_step, gpu_launches=int(gpu_launches),
               h2d=int(pinned.numel() * 4 + n * 8 + n * A * 8), d2h=int(n * A * 4 + n * 8), stats=stats,
               e2e_steps_ms=[dict(upload=a, search=b, readback=c) for a, b, c in e2e_parts[args.warmup:]])
    # ---- per-kernel profile (rank 0): eager steps of one search with CUDA events ------------------
    if rank == 0:
        planes, players, mcs = roots[-1]
        forest.set_roots_planes(planes, players, mcs)
        forest.begin_move(iters)
        adv, lv, per_step = [], [], {}
        eng.profile(True)
        eng.profile(False)                                            # zero the totals, timing off
        prev = {}
        for s in range(iters // vl + 1 if vl > 1 else iters):         # the whole search, every 8th step timed
            timed = s % 8 == 4
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            eng.profile_on(timed)
            a.record(stream)
            forest.advance()
            b.record(stream)
            eng.eval_leaves(forest, stream.cuda_stream)
            if timed:
                adv.append((a, b))
                lv.append(forest.counts()[0])
                cur = eng.profile_read()                              # per-step deltas: steps beyond 3x the median are dropped,
                for k, (ms, cnt) in cur.items():                      # so one preempted launch cannot skew a kernel's figure
                    pms, pcnt = prev.get(k, (0.0, 0))
                    if cnt > pcnt:
                        per_step.setdefault(k, []).append(((ms - pms) / (cnt - pcnt), cnt - pcnt))
                prev = cur
        eng.profile(False)
        leaves = sum(lv) / max(len(lv), 1)

        def robust_mean(xs):                                          # mean over the steps within 3x the median
            m = sorted(xs)[len(xs) // 2]
            keep = [x for x in xs if x <= 3.0 * m]
            return sum(keep) / len(keep)
        out["profile"] = {k: dict(ms=robust_mean([m for m, _ in v]), launches=sum(c for _, c in v)) for k, v in per_step.items()}
        out["profile"]["forest_advance"] = dict(ms=robust_mean([a.elapsed_time(b) for a, b in adv]), launches=len(adv))
        out["leaves_per_step"] = leaves
        out["proj_width"] = eng.proj_width()
        out["path_flags"] = eng.path_flags()
    out["clocks"] = clocks.summary(t_clock0, t_clock1)
    forest.close()
    eng.close()
    # ---- whole games through the public API (train.play_games): games/s, SURVEY 8d ------------------
    if rank == 0 and world == 1 and not args.no_selfplay:
        from alpha_zero_b200 import train as az_train
        np.random.seed(4321)
        az_train.play_games(G, model, min(n, 256), min(iters, 32), vl=vl, concurrent=min(n, 256))   # warm-up: engine + forest of the small size
        t0 = time.perf_counter()
        games = az_train.play_games(G, model, 2 * n, iters, vl=vl, dirichlet=True, concurrent=n)
        dt = time.perf_counter() - t0
        plies = sum(len(g[0]) for g in games)
        out["selfplay"] = dict(games=len(games), plies=plies, seconds=dt, games_per_s=len(games) / dt, positions_per_s=plies / dt,
                               simulations_per_s=plies * iters / dt, sims_per_move=iters, concurrent=n,
                               how="train.play_games: every game from the empty board to its end, Dirichlet noise on, moves "
                                   "sampled / argmax as the reference; 2 x concurrent games, a finished game's slot is refilled until all have "
                                   "started (the batch thins out only at the end)")
    if world > 1:
        dist.destroy_process_group()
    return out


def main():
    args = parse()
    wl = dict(WORKLOADS[args.workload])
    for k in ("trees", "iters", "vl"):
        if getattr(args, k):
            wl[k] = getattr(args, k)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    config = dict(workload=wl["name"], game=wl["game"], board=f"{wl['rows']}x{7 if wl['game'] == 'connect4' else wl['rows']}",
                  concurrent_games_per_gpu=wl["trees"], sims_per_move=wl["iters"], vl_batch=wl["vl"],
                  net="Net(D=512,patch=5,heads=8,depth=1) random-init, ai/nn.py / main.py:134",
                  l2="inputs larger than L2: forest node pools + activations >> 126 MB, fresh root positions every step",
                  parallelism=f"dp{world} (games sharded per GPU, no data-path collective)")
    base = dict(metric="mcts_simulations_per_second", unit="simulations/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                higher_is_better=True, scaling="weak", vs_baseline=None, dtype="bf16", data="synthetic", config=config)

    if args.impl == "reference":
        if rank != 0:
            return
        cb = run_cpu_baseline(wl, args.cpu_seconds, args.cpu_procs or min(os.cpu_count() or 1, 64))
        line = dict(base, impl="reference", value=cb["value"], ms_per_step=wl["trees"] * wl["iters"] * world / max(cb["value"], 1e-9) * 1e3,
                    cpu_baseline=cb, e2e=dict(value=cb["value"], unit="simulations/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                    gpu_launches=0)
        print(json.dumps(line))
        return

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb = run_cpu_baseline(wl, args.cpu_seconds, args.cpu_procs)
    r = run_b200(args, wl, rank, world)
    if rank != 0:
        return
    pk = peaks()
    value = r["sims_per_step"] * args.steps / (r["dev_ms"] / 1e3)
    e2e = r["sims_per_step"] * args.steps / r["e2e_s"]
    line = dict(base, impl="b200", value=value, ms_per_step=r["dev_ms"] / args.steps, clocks=r["clocks"],
                e2e=dict(value=e2e, unit="simulations/s", h2d_bytes_per_step=r["h2d"], d2h_bytes_per_step=r["d2h"],
                         steps_ms=r["e2e_steps_ms"]),
                gpu_launches=r["gpu_launches"], search_stats=r["stats"])
    pf = r.get("path_flags", 0)
    line["evaluator_path"] = dict(flags=pf, front_kernel="az_front2_kernel" if pf & 2 else ("az_front_kernel" if pf & 1 else "generic"),
                                  cls_row_gemm="cta_pair" if pf & 4 else "single_cta", layernorm_fold=bool(pf & 8), head_epilogue=bool(pf & 16))
    prof = r.get("profile", {})
    if prof:
        line["kernel_ms"] = {k: round(v["ms"], 4) for k, v in prof.items()}
        line["roofline"] = roofline(wl, r, prof, pk)
        line["rooflines_all"] = {k: rl for k, rl in ((k, roofline(wl, r, prof, pk, k)) for k in prof) if rl}
    if "selfplay" in r:
        line["selfplay"] = r["selfplay"]
    if cb:
        line["cpu_baseline"] = cb
    print(json.dumps(line))


if __name__ == "__main__":
    main()
