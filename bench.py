#!/usr/bin/env python
"""Benchmark of the MCTS hot path (BASELINE.json metric: MCTS simulations/s, 15x15, "hard").

    python bench.py --gpus N --steps K --warmup W            # the B200 engine
    python bench.py --impl reference --steps K --warmup W    # the CPU arm: the unmodified Python reference from
                                                             # baseline/_ref (one process per core), else the C port

A step = one search pass (gz_search) over one batch of G synthetic positions: S simulations per
position, each simulation = select + expand + policy-value forward of the new node + planner-guided
playout + backup.  Workload at N=1 = BASELINE config 3: hard preset, S=200,
c_puct=1.8, beta=0.2, planner_steps=5, G=4096 concurrent games.  With N>1 every rank runs its own
G games (weak scaling, no collective on the hot path; game ids are global so per-game random
streams do not depend on N).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "alphazero-gomoku_b200"))

# algorithmic FLOPs (2 x MACs) per unit, from the layer shapes
F_PV_CONV0 = 2 * 27 * 128 * 225
F_PV_TOWER = 4 * 2 * 1152 * 128 * 225                  # four 128->128 3x3 convolutions
F_PV_HEADS = 2 * (128 * 3 * 225 + 450 * 225 + 225 * 64 + 64)
F_GN_CONVS = 2 * 27 * 64 * 225 + 4 * (2 * 576 * 64 * 225 + 2 * 64 * 64 * 225)
F_GN_HEAD = 2 * (64 * 2 * 225 + 450 * 225)
F_DQN = 2 * (675 * 256 + 256 * 256 + 256 * 225)
STAGE_FLOPS = {"pv_conv0": ("pv_evals", F_PV_CONV0), "pv_tower": ("pv_evals", F_PV_TOWER),
               "pv_heads": ("pv_evals", F_PV_HEADS), "gn_convs": ("planner_steps", F_GN_CONVS),
               "gn_head": ("planner_steps", F_GN_HEAD), "dqn": ("planner_steps", F_DQN)}
# algorithmic HBM bytes per unit for the integer kernels (board in + result out)
STAGE_BYTES = {"tree": ("simulations", 80 + 100 + 48), "planner_score": ("planner_steps", 80 + 68),
               "planner_pick": ("planner_steps", 80 + 80 + 16 * 8 + 68), "rollout_tail": ("simulations", 80 + 80 + 12)}

PRESET = dict(sims=200, c_puct=1.8, beta=0.2, planner_steps=5, difficulty="hard")


_JSON_OUT = None


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on
    fd 1 under torchrun), so fd 1 is pointed at stderr for the whole run and the JSON line goes to a private
    duplicate of the original stdout."""
    global _JSON_OUT
    if _JSON_OUT is None:
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(line + "\n")
    out.flush()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sustained=float(p["bf16_tflops_sustained"]), source="measured")
    except Exception:
        return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler(threading.Thread):
    """Samples nvidia-smi SM clocks and throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4)
                          if len(r) > 2 + i and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def synthetic_weights():
    """Random-init weights of the three architectures (torch default init, seed 1234)."""
    import torch
    import bg_planner
    import neural_network
    torch.manual_seed(1234)
    pv = neural_network.AlphaZeroGomokuNet()
    gn, dq = bg_planner.GraphNet(), bg_planner.OpponentDQN()
    return pv.state_dict(), gn.state_dict(), dq.state_dict()


# ---------------------------------------------------------------------------------------- CPU arm
def cpu_sample(cells, to_move, count, seconds, sims_per_search, threads=None, fixed_searches=None):
    """Times the oracle port (oracle/gomoku_oracle.c: the reference's algorithm restated in C, pinned
    to the Python reference's goldens) on the host cores: every thread runs whole searches of
    `sims_per_search` simulations with the workload's preset on its share of the positions."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    pv, gn, dq = synthetic_weights()
    O = orc.Oracle(pv={k: v.numpy() for k, v in pv.items()}, graphnet={k: v.numpy() for k, v in gn.items()},
                   dqn={k: v.numpy() for k, v in dq.items()})
    threads = threads or os.cpu_count() or 1
    done = [0] * threads
    deadline = time.perf_counter() + seconds

    def work(t):
        i = t
        n = 0
        while (fixed_searches is None and time.perf_counter() < deadline) or \
                (fixed_searches is not None and n < fixed_searches):
            j = i % len(cells)
            b = orc.board_from_cells(cells[j], to_move[j], count[j])
            p = O.params(sims_per_search, PRESET["c_puct"], PRESET["beta"], PRESET["planner_steps"],
                         int(to_move[j]), PRESET["difficulty"], eval_nodes=1)
            r = O.search(b, p, seed=12345 + j, game=j)     # ctypes releases the GIL
            done[t] += sims_per_search
            i += threads
            n += 1
            del r

    t0 = time.perf_counter()
    ths = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    dt = time.perf_counter() - t0
    return sum(done) / dt, threads, sum(done), dt


def config_dict(G, n_gpus, extra=None):
    c = {"workload": "hard preset search: S=%d sims/position, c_puct=%.1f, beta=%.1f, planner_steps=%d, "
                     "G=%d concurrent 15x15 games per GPU (BASELINE config 3); synthetic mid-game "
                     "positions with %s stones" % (PRESET["sims"], PRESET["c_puct"], PRESET["beta"],
                                                   PRESET["planner_steps"], G, "6/12/24/48/96/150"),
         "games_per_gpu": G, "sims_per_position": PRESET["sims"], "parallelism": "games x%d" % n_gpus,
         "l2_policy": "per-step working set (activations, trees) exceeds the 126 MB L2; no explicit flush",
         "weights": "random init (torch default, seed 1234)"}
    if extra:
        c.update(extra)
    return c


REF_DIR = os.path.join(ROOT, "baseline", "_ref")
REF_FILES = ("gomoku_board.py", "ai_agent.py", "bg_planner.py", "neural_network.py")


def reference_available():
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in REF_FILES)


class _StopSearch(Exception):
    pass


def _ref_worker(w, nw, counter, ready, stop, sims_per_search):
    """One process = one core: the UNMODIFIED reference (baseline/_ref/{gomoku_board,ai_agent,bg_planner,
    neural_network}.py, copied there by __graft_entry__.build()) running AlphaZeroGomokuAI._mcts_search with the
    workload's preset on a spread of the workload positions.  Hygiene per SURVEY.md 8c: scratch cwd without ./models,
    time_limit = inf, one torch thread, params overridden through ai.params (training.py:240).  Nothing of the
    reference is modified: completed simulations are counted by wrapping the bound method _mcts_simulation."""
    import importlib.util
    import tempfile
    os.chdir(tempfile.mkdtemp(prefix="gz_ref_"))
    pkg = os.path.join(ROOT, "alphazero-gomoku_b200")
    spec = importlib.util.spec_from_file_location("gz_workload", os.path.join(pkg, "workload.py"))
    workload = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(workload)
    sys.path[:] = [REF_DIR] + [p for p in sys.path if os.path.abspath(p or ".") != pkg]
    import torch
    torch.set_num_threads(1)
    torch.manual_seed(1234)
    import ai_agent
    from gomoku_board import GomokuBoard
    assert os.path.dirname(os.path.abspath(ai_agent.__file__)) == REF_DIR
    cells, tm, mc = workload.make_positions(256, seed=2026)
    order = np.random.default_rng(7).permutation(256)        # every worker sees every stone count
    ais = {}
    for player in (1, 2):
        ai = ai_agent.AlphaZeroGomokuAI(player, "hard", device="cpu", beta=PRESET["beta"],
                                        planner_steps=PRESET["planner_steps"], time_limit=float("inf"))
        ai.params["num_simulations"] = sims_per_search
        ai.params["c_puct"] = PRESET["c_puct"]
        inner = ai._mcts_simulation

        def counted(root, _inner=inner):
            if stop.is_set():
                raise _StopSearch()
            _inner(root)
            with counter.get_lock():
                counter.value += 1
        ai._mcts_simulation = counted
        ais[player] = ai
    ready.set()
    k = w
    try:
        while not stop.is_set():
            i = int(order[k % 256])
            k += nw
            b = GomokuBoard()
            c = cells[i].reshape(15, 15)
            b.board = c.astype(b.board.dtype)
            blk = [(int(r), int(q), 1) for r, q in zip(*np.nonzero(c == 1))]
            wht = [(int(r), int(q), 2) for r, q in zip(*np.nonzero(c == 2))]
            b.move_history = [x for pair in zip(blk, wht) for x in pair] + blk[len(wht):]
            b.current_player = int(tm[i])
            ais[int(tm[i])]._mcts_search(b, b.get_valid_moves())
    except _StopSearch:
        pass


class ReferencePool:
    """os.cpu_count() worker processes running the Python reference; window(seconds) returns the simulations all of
    them completed in that wall-clock window."""

    def __init__(self, sims_per_search=2, workers=None):
        import multiprocessing as mp
        ctx = mp.get_context("spawn")
        self.n = workers or os.cpu_count() or 1
        self.stop = ctx.Event()
        self.counters = [ctx.Value("q", 0) for _ in range(self.n)]
        readies = [ctx.Event() for _ in range(self.n)]
        self.procs = [ctx.Process(target=_ref_worker, args=(w, self.n, self.counters[w], readies[w], self.stop,
                                                            sims_per_search), daemon=True) for w in range(self.n)]
        for p in self.procs:
            p.start()
        for r, p in zip(readies, self.procs):
            while not r.wait(1.0):
                if not p.is_alive():
                    raise RuntimeError("reference worker died during start-up")

    def total(self):
        return sum(c.value for c in self.counters)

    def window(self, seconds):
        n0, t0 = self.total(), time.perf_counter()
        time.sleep(seconds)
        return self.total() - n0, time.perf_counter() - t0

    def close(self):
        self.stop.set()
        for p in self.procs:
            p.join(timeout=20)
            if p.is_alive():
                p.terminate()


def run_reference(args):
    """CPU arm.  With baseline/_ref present: the reference's own Python code, one process per host core, each step a
    wall-clock window of --ref-step-seconds in which the workers keep running 2-simulation searches of the same
    preset over the workload positions (kind "reference"); the C port on all host threads is reported beside it as
    `optimised_cpu`.  Without baseline/_ref the port alone is timed (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import workload
    cells, tm, mc = workload.make_positions(256, seed=2026)
    sims_ref = 2
    threads = args.ref_workers or os.cpu_count() or 1
    extra = {}
    if reference_available() and not args.port_only:
        pool = ReferencePool(sims_ref, threads)
        try:
            for _ in range(args.warmup):
                pool.window(args.ref_step_seconds)
            total, t_all = 0, 0.0
            for _ in range(args.steps):
                n, dt = pool.window(args.ref_step_seconds)
                total += n
                t_all += dt
        finally:
            pool.close()
        kind = "reference"
        sample = ("%d processes (one per core, torch 1 thread each) running the unmodified reference's _mcts_search "
                  "(%d-simulation searches, time_limit=inf) over a fixed shuffle of 256 workload positions; a step = a "
                  "%.1f s wall-clock window, simulations counted as they complete" % (threads, sims_ref,
                                                                                     args.ref_step_seconds))
        rate, th, n_port, dt_port = cpu_sample(cells, tm, mc, min(args.cpu_seconds, 8.0), sims_ref, threads)
        extra["optimised_cpu"] = {"value": rate, "unit": "sims/s", "cores": th, "kind": "port",
                                  "sample": "C restatement (oracle/), %d sims in %.1f s" % (n_port, dt_port)}
    else:
        for _ in range(args.warmup):
            cpu_sample(cells, tm, mc, 0, sims_ref, threads, fixed_searches=1)
        total, t_all = 0, 0.0
        for _ in range(args.steps):
            rate, th, n, dt = cpu_sample(cells, tm, mc, 0, sims_ref, threads, fixed_searches=2)
            total += n
            t_all += dt
        kind = "port"
        sample = ("baseline/_ref absent: C port; %d threads x 2 searches x %d sims per step on the first positions of "
                  "the workload" % (threads, sims_ref))
    value = total / max(t_all, 1e-9)
    line = {"impl": "reference", "metric": "mcts_sims_per_sec", "value": value, "unit": "sims/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_all / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args.games, args.gpus),
            "cpu_baseline": {"value": value, "unit": "sims/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "sims/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    line.update(extra)
    emit(json.dumps(line))


def ncu_traffic(stage):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the stage's main kernel, from the committed
    `ncu --set full` capture of this round (profiles/ncu_traffic.json: written by profiles/ncu_key_metrics.py from the
    .ncu-rep of the same bench command); None when no capture is on record for the stage."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(stage, {}).get("bytes_per_launch")
    except Exception:
        return None


def steady_selfplay(eng, G, rank, warm_plies, window_plies, torch):
    """BASELINE's second metric inside the same run: hard-preset self-play with game recycling.  Warm start: the G
    games are first spread over all ages with `warm_plies` cheap plies (8 simulations per move, same engine, same
    rules) so that slots finish and recycle at different times; then the budget goes back to S = 200 and
    `window_plies` plies (>= one mean game) are timed.  Returns (milliseconds, stats delta, counters delta)."""
    import engine
    p = engine.make_params(8, PRESET["c_puct"], beta=PRESET["beta"], planner_steps=PRESET["planner_steps"],
                           difficulty=PRESET["difficulty"], deterministic=False, seed=4242, game_base=rank * G)
    eng.selfplay_begin(G, p, 0.01)        # hard preset: exploration 0.01 (ai_agent.py:84-90)
    for _ in range(warm_plies):
        eng.selfplay_step()
    eng.selfplay_drain()
    eng.selfplay_set_simulations(PRESET["sims"])
    for _ in range(3):
        eng.selfplay_step()
    torch.cuda.synchronize()
    st0, c0 = eng.selfplay_stats(), eng.counters()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for k in range(window_plies):
        eng.selfplay_step()
        if k % 32 == 31:
            eng.selfplay_drain()
    ev1.record()
    torch.cuda.synchronize()
    st1, c1 = eng.selfplay_stats(), eng.counters()
    return ev0.elapsed_time(ev1), {k: st1[k] - st0[k] for k in st1}, {k: c1[k] - c0[k] for k in c1}


def run_engine(args):
    import torch
    import engine
    import workload
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    G, S = args.games, PRESET["sims"]
    S_CODE = 400                                   # the code's own "hard" preset (ai_agent.py:84-90)
    eng = engine.Engine(max_games=G, max_simulations=S_CODE if args.s400_steps else S, device=local,
                        store_priors=True, precision=1 if args.precision == "bf16" else 0, lanes=args.lanes)
    pv, gn, dq = synthetic_weights()
    eng.set_pv_weights(pv)
    eng.set_planner_weights(gn, dq)
    cells, tm, mc = workload.make_positions(G, seed=2026 + rank)
    boards = eng.boards_from_cells(cells, tm, mc)
    ai = tm.copy()

    def params(step, sims=S):
        return engine.make_params(sims, PRESET["c_puct"], beta=PRESET["beta"],
                                  planner_steps=PRESET["planner_steps"], difficulty=PRESET["difficulty"],
                                  deterministic=False, seed=0xB200 + step, game_base=rank * G)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def reduce_max(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def reduce_sum(x):
        t = torch.tensor([float(x)], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.item()

    # ---- the timed region: K searches, inputs resident in HBM, no instrumentation inside
    for w in range(args.warmup):
        eng.search(boards, ai, params(1000 + w))
    barrier()
    eng.reset_counters()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for k in range(args.steps):
        eng.search(boards, ai, params(k))
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.finish()
    cnt = eng.counters()
    ms_max, sims_total = reduce_max(ms), reduce_sum(cnt["simulations"])
    value = sims_total / (ms_max / 1e3)

    # ---- end to end through the host-buffer entry point (pinned host memory, copies timed)
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()  # noqa: E731
    h_cells, h_tm, h_mc, h_ai = pin(cells), pin(tm), pin(mc), pin(ai)
    out = (pin(np.empty((G, 225), np.int32)), pin(np.empty((G, 225), np.float64)), pin(np.empty(G, np.int32)))
    eng.search_host(h_cells, h_tm, h_mc, h_ai, params(2000), out)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        eng.search_host(h_cells, h_tm, h_mc, h_ai, params(3000 + k), out)
    torch.cuda.synchronize()
    e2e_s = reduce_max(time.perf_counter() - t0)
    e2e = world * G * S * args.steps / e2e_s
    h2d = G * (225 + 1 + 1 + 2)
    d2h = G * (225 * 4 + 225 * 8 + 4)

    # ---- one extra step with stage events on (every kernel on one stream): where the time goes, and the roofline of
    #      the dominant kernel.  The ~19 k event records it needs stay out of the timed region above.
    eng.set_lanes(1)
    eng.reset_counters()
    eng.set_profiling(True)
    eng.stage_times()
    eng.search(boards, ai, params(5000))
    stages = eng.stage_times()
    cnt_stage = eng.counters()
    eng.set_profiling(False)
    eng.set_lanes(args.lanes)

    # ---- the code's own hard preset (S = 400) on the same positions
    s400 = None
    if args.s400_steps:
        eng.search(boards, ai, params(6000, S_CODE))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for k in range(args.s400_steps):
            eng.search(boards, ai, params(6100 + k, S_CODE))
        e1.record()
        barrier()
        ms400 = reduce_max(e0.elapsed_time(e1))
        s400 = {"value": world * G * S_CODE * args.s400_steps / (ms400 / 1e3), "unit": "sims/s",
                "sims_per_position": S_CODE, "steps": args.s400_steps, "ms_per_step": ms400 / args.s400_steps,
                "note": "ai_agent.py:84-90 preset (400 simulations, c_puct 1.8); same positions, device-resident"}

    # ---- BASELINE's second metric: self-play games/hour, steady state
    sp = None
    if args.selfplay_plies:
        barrier()
        sp_ms, sp_st, sp_cnt = steady_selfplay(eng, G, rank, args.selfplay_warm, args.selfplay_plies, torch)
        t_sp = reduce_max(sp_ms) / 1e3
        games, plies = reduce_sum(sp_st["games_finished"]), reduce_sum(sp_st["plies_played"])
        sp = {"value": games / t_sp * 3600.0, "unit": "games/h", "window_plies": args.selfplay_plies,
              "warm_start": "%d plies at 8 simulations per move to spread the games over all ages, then 3 plies at "
                            "S=200 before the window" % args.selfplay_warm,
              "games_finished": games, "plies_played": plies, "seconds": t_sp,
              "mean_plies_per_finished_game": plies / games if games else None,
              "sims_per_sec": reduce_sum(sp_cnt["simulations"]) / t_sp,
              "records_dropped": reduce_sum(sp_st["records_dropped"]),
              "config": "hard preset self-play from empty boards with game recycling: S=200, c_puct=1.8, beta=0.2, "
                        "planner_steps=5, epsilon=0.01, %d concurrent games per GPU" % G}

    if rank == 0:
        peaks = load_peaks()
        total_ms = sum(v[0] for v in stages.values())
        dom = max(stages, key=lambda k: stages[k][0])
        d_ms, d_launch = stages[dom]
        share = {k: round(v[0] / total_ms, 4) for k, v in stages.items() if v[1]}
        if dom in STAGE_FLOPS:
            unit, f = STAGE_FLOPS[dom]
            achieved = cnt_stage[unit] * f / (d_ms / 1e3) / 1e12
            roof = {"bound": "tensor", "achieved": achieved, "peak": peaks["tf_sustained"],
                    "unit": "TFLOP/s", "frac": achieved / peaks["tf_sustained"], "traffic": ncu_traffic(dom)}
        else:
            unit, by = STAGE_BYTES.get(dom, ("simulations", 0))
            achieved = cnt_stage[unit] * by / (d_ms / 1e3) / 1e9
            roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm"], "unit": "GB/s",
                    "frac": achieved / peaks["hbm"], "traffic": ncu_traffic(dom)}
        roof.update({"kernel": dom, "launches": d_launch, "avg_launch_ms": d_ms / max(d_launch, 1),
                     "share_of_step": round(d_ms / total_ms, 4), "stage_share": share,
                     "stage_ms": {k: round(v[0], 2) for k, v in stages.items() if v[1]},
                     "measured": "CUDA events around every kernel group of ONE extra step after the timed region "
                                 "(gz_set_profiling; every kernel on the launching stream)",
                     "peak_source": peaks["source"] + (" (sustained bf16 cuBLAS)" if roof["bound"] == "tensor" else " (copy)")})
        f_sim = (cnt["pv_evals"] * (F_PV_CONV0 + F_PV_TOWER + F_PV_HEADS) +
                 cnt["planner_steps"] * (F_GN_CONVS + F_GN_HEAD + F_DQN)) / max(cnt["simulations"], 1)
        line = {"metric": "mcts_sims_per_sec", "value": value, "unit": "sims/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.precision,
                "data": "synthetic", "config": config_dict(G, world),
                "clocks": clocks,
                "e2e": {"value": e2e, "unit": "sims/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
                "gpu_launches": cnt["kernel_launches"],
                "roofline": roof,
                "flops_per_sim": f_sim,
                "tensor_frac_whole_step": value * f_sim / 1e12 / peaks["tf_sustained"] / world,
                "counters": cnt}
        if sp:
            line["selfplay_games_per_hour"] = sp
        if s400:
            line["code_preset_s400"] = s400
        if world == 1:
            line.update(cpu_legs(args, cells, tm, mc))
        emit(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def cpu_legs(args, cells, tm, mc):
    """cpu_baseline of the engine line (rank 0, N = 1): the Python reference from baseline/_ref on every host core for
    --cpu-seconds when it is there (kind "reference"), the C port beside it as optimised_cpu; else the port."""
    out = {}
    threads = args.ref_workers or os.cpu_count() or 1
    port_rate, cores, n_cpu, dt_cpu = cpu_sample(cells[:256], tm[:256], mc[:256],
                                                 args.cpu_seconds if not reference_available() else min(args.cpu_seconds, 8.0),
                                                 2, threads)
    port = {"value": port_rate, "unit": "sims/s", "cores": cores, "kind": "port",
            "sample": "%d sims in %.1f s: %d threads running 2-sim searches of the same preset on the first 256 "
                      "workload positions (C restatement, oracle/)" % (n_cpu, dt_cpu, cores)}
    if reference_available() and not args.port_only:
        try:
            pool = ReferencePool(2, threads)
            try:
                pool.window(2.0)
                n, dt = pool.window(args.cpu_seconds)
            finally:
                pool.close()
            out["cpu_baseline"] = {"value": n / dt, "unit": "sims/s", "cores": threads, "kind": "reference",
                                   "sample": "%d sims in %.1f s: %d processes running the unmodified Python reference's "
                                             "_mcts_search (2-sim searches, same preset) over a shuffle of 256 workload "
                                             "positions" % (n, dt, threads)}
            out["optimised_cpu"] = port
            return out
        except Exception as exc:          # the arm must not take the GPU line down with it
            port["sample"] += "; Python reference arm failed: %r" % (exc,)
    out["cpu_baseline"] = port
    return out


def selfplay_session(eng, G, rank, seed=2026):
    import engine
    p = engine.make_params(PRESET["sims"], PRESET["c_puct"], beta=PRESET["beta"],
                           planner_steps=PRESET["planner_steps"], difficulty=PRESET["difficulty"],
                           deterministic=False, seed=seed, game_base=rank * G)
    eng.selfplay_begin(G, p, 0.01)        # hard preset: exploration 0.01 (ai_agent.py:84-90)


def run_selfplay(args):
    """Second metric of BASELINE.json: self-play games/hour, hard preset (S=200, planner_steps=5,
    epsilon 0.01), G concurrent games per GPU with recycling, every ply one full gz_search.
    Timed: from gz_selfplay_begin until every rank has finished --min-games games (CUDA events, max
    over ranks); value = games finished by all ranks / that time."""
    import torch
    import engine
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    G = args.games
    eng = engine.Engine(max_games=G, max_simulations=PRESET["sims"], device=local, store_priors=True,
                        precision=1 if args.precision == "bf16" else 0, lanes=args.lanes)
    pv, gn, dq = synthetic_weights()
    eng.set_pv_weights(pv)
    eng.set_planner_weights(gn, dq)
    selfplay_session(eng, G, rank, seed=1)
    for _ in range(args.warmup):
        eng.selfplay_step()
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    eng.reset_counters()
    selfplay_session(eng, G, rank)
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    steps = 0
    while True:
        eng.selfplay_step()
        steps += 1
        if steps % 8 == 0 or steps >= args.max_plies:
            st = eng.selfplay_stats()
            if st["records_ready"] > G * 128:
                eng.selfplay_drain()
            if st["games_finished"] >= args.min_games or steps >= args.max_plies:
                break
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.finish()
    st = eng.selfplay_stats()
    cnt = eng.counters()
    vals = torch.tensor([ms, st["games_finished"], st["plies_played"], cnt["simulations"], st["black_wins"],
                         st["white_wins"], st["draws"]], dtype=torch.float64, device="cuda")
    mx = vals.clone()
    if dist is not None:
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(vals, op=dist.ReduceOp.SUM)
    if rank == 0:
        t = mx[0].item() / 1e3
        games, plies, sims = vals[1].item(), vals[2].item(), vals[3].item()
        emit(json.dumps({
            "metric": "selfplay_games_per_hour", "value": games / t * 3600.0, "unit": "games/h", "n_gpus": world,
            "steps": steps, "warmup": args.warmup, "ms_per_step": mx[0].item() / steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
            "config": config_dict(G, world, {"workload": "hard preset self-play from empty boards with game "
                                             "recycling: S=200, c_puct=1.8, beta=0.2, planner_steps=5, epsilon=0.01, "
                                             "G=%d concurrent games per GPU" % G}),
            "clocks": clocks, "games_finished": games, "plies_played": plies,
            "mean_plies_per_finished_game": (plies / games) if games else None,
            "sims_per_sec": sims / t, "plies_per_sec": plies / t,
            "outcomes": {"black": vals[4].item(), "white": vals[5].item(), "draw": vals[6].item()},
            "gpu_launches": cnt["kernel_launches"], "seconds": t}))
    if dist is not None:
        dist.destroy_process_group()


def run_iteration(args):
    """BASELINE config 5: one training.py iteration (training.py:401-508) -- hard-preset self-play on every GPU until
    --min-games games per rank are finished, NCCL all-gather of the packed replay, device dataset build (35 % x 8
    symmetries), two epochs of the CE+MSE Adam update (batch 128 x ranks; every rank runs the reference's single
    trainer on the gathered set), weight broadcast, evaluation arena.  Reports games/h of the WHOLE iteration and the
    wall time of every phase (max over ranks)."""
    import torch
    import neural_network
    import training
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.manual_seed(1234)
    gm = neural_network.GomokuModel(device=str(dev))
    opt = torch.optim.Adam(gm.model.parameters(), lr=8e-4, weight_decay=1e-5, capturable=not args.eager_update)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=2, gamma=0.85)
    kw = dict(games_per_iteration=args.min_games * world, concurrent=args.games, difficulty="hard", beta=0.2,
              planner_steps=5, num_simulations=PRESET["sims"], c_puct=PRESET["c_puct"], batch_size=128 * world,
              eval_games=args.eval_games, rank=rank, world=world, precision=args.precision)
    if args.warmup:   # a short untimed iteration: allocations, cuDNN autotune, NCCL rings
        training.run_iteration(gm, opt, None, dev, seed=1, **dict(kw, games_per_iteration=world, num_simulations=8,
                                                                  eval_games=0))
    if dist is not None:
        dist.barrier()
    sampler = ClockSampler(local)
    sampler.start()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = training.run_iteration(gm, opt, sched, dev, seed=2, **kw)
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    clocks = sampler.finish()
    ph = out["phases"]
    vals = torch.tensor([t, ph["selfplay_s"], ph["gather_s"], ph["dataset_s"], ph["update_s"], ph["eval_s"]],
                        dtype=torch.float64, device="cuda")
    games = torch.tensor([float(out["games"])], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(games, op=dist.ReduceOp.SUM)
    if rank == 0:
        tt = vals[0].item()
        emit(json.dumps({
            "metric": "training_iteration_games_per_hour", "value": games.item() / tt * 3600.0, "unit": "games/h",
            "n_gpus": world, "steps": 1, "warmup": int(bool(args.warmup)), "ms_per_step": tt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": args.precision,
            "data": "synthetic",
            "config": config_dict(args.games, world, {"workload": "training.py full iteration (BASELINE config 5): hard "
                                  "preset self-play (%d games per GPU, %d concurrent) + replay all-gather + dataset "
                                  "build + 2 epochs CE+MSE Adam + %d evaluation games" %
                                  (args.min_games, args.games, args.eval_games)}),
            "clocks": clocks, "games": games.item(), "records": out["records"], "samples": out["samples"],
            "train_loss": out.get("train_loss"), "val_loss": out.get("val_loss"), "eval": out.get("eval"),
            "phases_s": {k: round(v, 4) for k, v in zip(("total", "selfplay", "gather", "dataset", "update", "eval"),
                                                        vals.tolist())}}))
    if dist is not None:
        dist.destroy_process_group()


def run_latency(args):
    """BASELINE config 1 through the drop-in classes: ONE self-play game, easy preset as the README quotes it (50
    simulations, c_puct 1.2, set through ai.params like training.py:240), beta 0.2, planner_steps 5, two
    AlphaZeroGomokuAI objects sharing one GomokuModel (training.py:403-408), G = 1 per get_move call: the interactive
    call pattern of game_ui.py:178 / training.py:173.  Reports milliseconds per searched move and per game; every
    get_move includes its host<->device copies.  --games-latency games are played, the first is warm-up."""
    import random
    import torch
    from ai_agent import AlphaZeroGomokuAI
    from gomoku_board import GomokuBoard
    from neural_network import GomokuModel
    torch.manual_seed(1234)
    random.seed(1234)
    shared = GomokuModel(device="cuda")
    ais = {}
    for player in (1, 2):
        ai = AlphaZeroGomokuAI(player, "easy", device="cuda", beta=0.2, planner_steps=5, time_limit=float("inf"),
                               precision=args.precision, seed=99 + player)
        ai.model = shared
        ai.params["num_simulations"] = 50
        ai.params["c_puct"] = 1.2
        ais[player] = ai
    sampler = ClockSampler(0)
    games = []
    for gi in range(1 + args.games_latency):
        if gi == 1:
            sampler.start()
        b = GomokuBoard()
        per_move, t_game = [], time.perf_counter()
        while not b.game_over:
            ai = ais[b.current_player]
            searched = b.get_move_count() >= 6
            t0 = time.perf_counter()
            mv = ai.get_move(b)
            dt = time.perf_counter() - t0
            if mv is None:
                break
            if searched:
                per_move.append(dt * 1e3)
            b.make_move(*mv)
        games.append(dict(plies=b.get_move_count(), seconds=time.perf_counter() - t_game, ms=per_move,
                          winner=b.winner))
    clocks = sampler.finish()
    timed = games[1:]
    ms = sorted(x for g in timed for x in g["ms"])
    launches = sum(ai._engine.counters()["kernel_launches"] for ai in ais.values() if ai._engine)
    med = ms[len(ms) // 2]
    emit(json.dumps({
        "metric": "get_move_latency_ms", "value": med, "unit": "ms per searched move (median)", "n_gpus": 1,
        "steps": len(ms), "warmup": 1, "ms_per_step": med, "higher_is_better": False, "scaling": "weak",
        "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": "BASELINE config 1: easy preset (50 sims, c_puct 1.2, beta 0.2, planner_steps 5, epsilon "
                               "0.2), one self-play game at a time through AlphaZeroGomokuAI.get_move, G = 1",
                   "weights": "random init (torch default, seed 1234)"},
        "clocks": clocks, "mean_ms": sum(ms) / len(ms), "p90_ms": ms[int(0.9 * (len(ms) - 1))], "max_ms": ms[-1],
        "sims_per_sec_g1": 50.0 / (sum(ms) / len(ms) / 1e3),
        "games": [dict(plies=g["plies"], seconds=round(g["seconds"], 4), winner=g["winner"]) for g in timed],
        "seconds_per_game": sum(g["seconds"] for g in timed) / len(timed),
        "gpu_launches": launches,
        "reference": "the Python reference needs ~9.5 s per move as shipped (time_limit 5 s: 2 simulations complete) "
                     "and ~400 s for 50 simulations on an early-game board (BASELINE.md section 2)"}))


def synthetic_records(n, seed):
    """n packed replay records (72 B) of random mid-game positions, vectorised: stone k of a random permutation
    of the cells goes to colour k & 1, the record's move is the next cell of the permutation."""
    import engine
    rng = np.random.default_rng(seed)
    rec = np.zeros(n, dtype=engine.REPLAY_DTYPE)
    order = np.argsort(rng.random((n, 225)), axis=1).astype(np.int64)
    k = rng.integers(0, 150, n)
    rank = np.empty_like(order)
    np.put_along_axis(rank, order, np.broadcast_to(np.arange(225), (n, 225)), axis=1)    # rank[i, cell]
    bitpos = 16 * (np.arange(225) // 15) + np.arange(225) % 15
    for colour in (0, 1):
        on = (rank < k[:, None]) & ((rank & 1) == colour)
        bits = np.zeros((n, 256), np.uint8)
        bits[:, bitpos] = on
        rec["stones"][:, colour] = np.packbits(bits, axis=1, bitorder="little").view("<u4").reshape(n, 8)
    rec["move"] = order[np.arange(n), k]
    rec["mover"] = 1 + (k & 1)
    rec["outcome"] = rng.integers(-1, 2, n)
    rec["ply"] = k
    rec["game_id"] = np.arange(n)
    return rec.view(np.uint8).reshape(n, 72)


def run_dataset(args):
    """GomokuSelfPlayDataset build (every record + 8 symmetries of 35 % of them, the ratio
    training.py:461 uses) from --records packed replay records per GPU.  A step decodes the whole dataset
    (n + 8 * int(0.35 n) samples of 2712 B) into device tensors; records are resident in HBM for `value`,
    come from pinned host memory for `e2e` (labels read back, planes stay on the device for the optimiser)."""
    import random
    import torch
    import engine
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    n = args.records
    eng = engine.Engine(max_games=8, max_simulations=4, device=local)
    rec_h = torch.from_numpy(synthetic_records(n, 99 + rank)).pin_memory()
    random.seed(1234 + rank)
    chosen = np.asarray(random.sample(range(n), k=max(1, int(n * 0.35))), np.int32)
    count = n + 8 * len(chosen)
    rec_d, ch_d = rec_h.cuda(), torch.from_numpy(chosen).cuda()
    out = None

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        out = eng.dataset_build(rec_d, ch_d, reference_exact=True)
    barrier()
    eng.reset_counters()
    sampler = ClockSampler(local)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        del out
        out = eng.dataset_build(rec_d, ch_d, reference_exact=True)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.finish()
    launches = eng.counters()["kernel_launches"]
    ch_h = torch.from_numpy(chosen).pin_memory()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        del out
        out = eng.dataset_build(rec_h.cuda(non_blocking=True), ch_h.cuda(non_blocking=True), reference_exact=True)
        labels = (out[1].cpu(), out[2].cpu())
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        peaks = load_peaks()
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle as orc
        import selfplay
        ns = min(n, 2048)
        x, mv, _m, oc = selfplay.records_to_samples(rec_h.numpy()[:ns])
        t1 = time.perf_counter()
        reps = 0
        while time.perf_counter() - t1 < min(args.cpu_seconds, 10.0):
            ox, _, _ = orc.dataset_samples(x, mv, oc, list(range(0, ns, 3)), reference_exact=True)
            reps += 1
        cpu_rate = reps * len(ox) / (time.perf_counter() - t1)
        per_sample = 675 * 4 + 8 + 4            # written
        bytes_step = count * per_sample + n * 72 + len(chosen) * 4
        gbs = bytes_step * args.steps / (t[0].item() / 1e3) / 1e9
        emit(json.dumps({
            "metric": "dataset_samples_per_sec", "value": world * count * args.steps / (t[0].item() / 1e3),
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": t[0].item() / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 bitboards -> f32", "data": "synthetic",
            "config": {"workload": "GomokuSelfPlayDataset build: %d records per GPU, augment_ratio 0.35 -> %d samples "
                                   "(training.py:104-134), reference_exact index transform" % (n, count),
                       "records_per_gpu": n, "samples_per_gpu": count,
                       "l2_policy": "output (%.1f GB per step) exceeds the 126 MB L2" % (count * per_sample / 1e9)},
            "clocks": clocks,
            "e2e": {"value": world * count * args.steps / (t[1].item() / 1e3), "unit": "samples/s",
                    "h2d_bytes_per_step": n * 72 + len(chosen) * 4, "d2h_bytes_per_step": count * 12},
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": peaks["hbm"], "unit": "GB/s",
                         "frac": gbs / peaks["hbm"], "traffic": None, "kernel": "k_dataset_build",
                         "algorithmic_bytes_per_sample": per_sample, "peak_source": peaks["source"] + " (copy)"},
            "cpu_baseline": {"value": cpu_rate, "unit": "samples/s", "cores": 1, "kind": "port",
                             "sample": "numpy restatement (oracle.dataset_samples) on %d records, %d passes" % (ns, reps)}}))
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--games", type=int, default=4096)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--ref-step-seconds", type=float, default=4.0,
                    help="--impl reference: wall-clock window of one step of the Python reference arm")
    ap.add_argument("--ref-workers", type=int, default=0, help="CPU legs: worker processes / threads (0 = all cores)")
    ap.add_argument("--port-only", action="store_true", help="CPU legs: time the C port even if baseline/_ref exists")
    ap.add_argument("--lanes", type=int, default=0,
                    help="independent game groups advanced on separate CUDA streams (0 = engine default)")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "f32"],
                    help="network arithmetic: bf16 tcgen05 (timed mode) or fp32 FFMA (parity mode)")
    ap.add_argument("--mode", default="search", choices=["search", "selfplay", "dataset", "iteration", "latency"],
                    help="search: sims/s on fixed positions (the headline); selfplay: games/hour; "
                         "dataset: training samples/s of the dataset build (SURVEY 8f row 2)")
    ap.add_argument("--records", type=int, default=1 << 20, help="dataset mode: replay records per GPU")
    ap.add_argument("--selfplay-plies", type=int, default=80,
                    help="search mode: plies of the steady-state self-play window that gives games/hour (0 = skip)")
    ap.add_argument("--selfplay-warm", type=int, default=160,
                    help="search mode: cheap 8-simulation plies that spread the games over all ages before the window")
    ap.add_argument("--s400-steps", type=int, default=2, help="search mode: extra steps at the code preset S=400 (0 = skip)")
    ap.add_argument("--games-latency", type=int, default=3, help="latency mode: timed games (one more is played first)")
    ap.add_argument("--min-games", type=int, default=4096, help="selfplay mode: games to finish per rank")
    ap.add_argument("--max-plies", type=int, default=2000)
    ap.add_argument("--eval-games", type=int, default=64, help="iteration mode: evaluation arena games")
    ap.add_argument("--eager-update", action="store_true",
                    help="iteration mode: run the update eagerly instead of replaying one captured CUDA graph per batch")
    args = ap.parse_args()
    claim_stdout()
    if args.mode == "selfplay" and args.impl != "reference":
        return run_selfplay(args)
    if args.mode == "iteration" and args.impl != "reference":
        return run_iteration(args)
    if args.mode == "dataset" and args.impl != "reference":
        return run_dataset(args)
    if args.mode == "latency" and args.impl != "reference":
        return run_latency(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_engine(args)


if __name__ == "__main__":
    main()
