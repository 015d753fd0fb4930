"""Do the integer kernels co-run with the persistent tensor kernels when they come from another stream?  Two engine
handles on two CUDA streams: A loops the planner networks (embed + 4 x k_gn_block + head/DQN) or the policy-value
tower, B loops an integer kernel family (rollout tail / planner score via gz_planner_move's first kernel is not
separable, so: gz_rollout with planner_steps=0, gz_board_ops, gz_rollout_policy).  Reports each alone and both together."""
import os, sys, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path[:0] = [os.path.join(ROOT, "alphazero-gomoku_b200"), ROOT]
import torch
import engine, workload, bench
G = 2048
pv, gn, dq = bench.synthetic_weights()
cells, tm, mc = workload.make_positions(G, seed=2026)
engs = []
for i in range(2):
    e = engine.Engine(max_games=G, max_simulations=8, precision=1)
    e.set_pv_weights(pv); e.set_planner_weights(gn, dq)
    engs.append(e)
sA, sB = torch.cuda.Stream(), torch.cuda.Stream(priority=-1)
boards = [e.boards_from_cells(cells, tm, mc) for e in engs]
p0 = engine.make_params(1, 1.8, planner_steps=0, difficulty="hard", deterministic=False, seed=3)
torch.cuda.synchronize()

def tensor_work(kind, n):
    with torch.cuda.stream(sA):
        for _ in range(n):
            if kind == "gn":
                engs[0].planner_nets_forward(boards[0])
            else:
                engs[0].pv_forward(boards[0])

def int_work(kind, n):
    with torch.cuda.stream(sB):
        for _ in range(n):
            if kind == "rollout":
                engs[1].rollout(boards[1], tm, p0)
            elif kind == "ops":
                engs[1].board_ops(boards[1])
            else:
                engs[1].rollout_policy(boards[1], p0)

def timed(f):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    f()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) * 1e3

for tk, nt in (("gn", 200), ("pv", 100)):
    for ik, ni in (("rollout", 400), ("ops", 1500), ("policy", 1500)):
        tensor_work(tk, 5); int_work(ik, 5)
        a = timed(lambda: tensor_work(tk, nt))
        b = timed(lambda: int_work(ik, ni))
        both = timed(lambda: (tensor_work(tk, nt), int_work(ik, ni)))
        print("%s x%d alone %.1f ms | %s x%d alone %.1f ms | together %.1f ms (serial sum %.1f, hidden %.0f %% of the shorter)"
              % (tk, nt, a, ik, ni, b, both, a + b, 100 * (a + b - both) / min(a, b)), flush=True)
