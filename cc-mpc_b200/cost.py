"""Closed-loop cost aggregation: twin of dataForCost_ref.py / montecarlo_cost_ref.py.

``cost_final`` (dataForCost_ref.py:198-232) for a BATCH of episodes runs in one
kernel launch (ccmpc_episode_cost); the pickle scan / episode bookkeeping of
montecarlo_cost_ref.py:10-93 stays host Python (it is file I/O).
"""
from __future__ import annotations

import os
import pickle
import re
from typing import List, Sequence

import numpy as np
import torch

from . import _lib, ops

GOAL = (167.174698, -81.759842)      # dataForCost_ref.py:209


def episode_costs(X_last, U, lengths, goal=GOAL):
    """Batched cost_final.  ``X_last`` [E,2] final stored position of each episode,
    ``U`` [E,Hmax,2] controls (padded), ``lengths`` [E] int32 = H per episode.
    Returns (cost [E], costU [E]) device tensors; H < 2 yields NaN (the reference raises)."""
    lib = _lib.load()
    ops._chk(X_last, torch.float64, "X_last", (2,))
    ops._chk(U, torch.float64, "U", (2,))
    ops._chk(lengths, torch.int32, "lengths")
    E, Hmax = U.shape[0], U.shape[1]
    cost = torch.empty((E,), dtype=torch.float64, device=U.device)
    costU = torch.empty((E,), dtype=torch.float64, device=U.device)
    _lib.check(lib.ccmpc_episode_cost(ops._p(X_last), ops._p(U), ops._p(lengths), E, Hmax, float(goal[0]),
                                      float(goal[1]), ops._p(cost), ops._p(costU), ops._stream()),
               "ccmpc_episode_cost")
    return cost, costU


def cost_final(X_star: Sequence, U_star: Sequence, goal=GOAL):
    """dataForCost_ref.py:198-232 for one episode.  Returns (cost, costU) floats."""
    H = len(U_star)
    if H < 2:
        raise UnboundLocalError("local variable 'costU' referenced before assignment")   # reference behaviour
    dev = torch.device("cuda", torch.cuda.current_device())
    xl = torch.tensor([[float(X_star[-1][0]), float(X_star[-1][1])]], dtype=torch.float64, device=dev)
    U = torch.tensor(np.asarray([np.asarray(u, dtype=np.float64).reshape(2) for u in U_star])[None],
                     dtype=torch.float64, device=dev)
    cost, costU = episode_costs(xl, U, torch.tensor([H], dtype=torch.int32, device=dev), goal)
    return float(cost.item()), float(costU.item())


class getData_singleEpisode:
    """dataForCost_ref.py:13-88: loads the ``agent{id}_frame{f}_cov`` pickles of one episode."""

    def __init__(self, Agent, frame_id, numFrames, root="out/data/"):
        self.Agent = Agent
        self.frame_id = frame_id
        self.numFrames = numFrames
        self.T = 8          # control horizon (:18)
        self.path = ""
        self.root = root

    def frame_list(self):
        return [self.frame_id + i * 10 for i in range(self.numFrames)]     # :21-25

    def get_data(self):
        """-> (avrProcessTime, avrSolveTime, avrSolveTimeOV, worstTime, X_star, U_star, goal, x_init, timeout)
        (:28-88).  Entries of infeasible frames stay ``None`` as in the reference."""
        nF = self.numFrames
        loaded = [None] * nF
        process_time, solve_time, X_star, U_star, goal, x_init, timeout, ONotZero = ([None] * nF for _ in range(8))
        for t, frame in enumerate(self.frame_list()):
            path = os.path.join(self.root + self.path, f"agent{self.Agent}_frame{frame}_cov")
            with open(path, "rb") as f:
                loaded[t] = pickle.load(f)
            d = loaded[t]
            if d["infeasible"] == False:  # noqa: E712  (reference comparison, :56)
                process_time[t] = d["process_time"]; solve_time[t] = d["solve_time"]
                X_star[t] = d["X_star"][0]; U_star[t] = d["U_star"][0]
                goal[t] = d["goal"]; x_init[t] = d["x_init"]; timeout[t] = d["timeout"]; ONotZero[t] = d["MeanCov"]
        processTime, solveTime, solveTimeOV = [], [], []
        worstTime = 0
        for t in range(nF):
            if loaded[t]["infeasible"] == False and ONotZero[t]:  # noqa: E712
                processTime.append(process_time[t]); solveTime.append(solve_time[t])
                worstTime = np.max([worstTime, solve_time[t]])
                solveTimeOV.append(solve_time[t])
        if processTime:
            avrP, avrS, avrSOV = (sum(processTime) / len(processTime), sum(solveTime) / len(solveTime),
                                  sum(solveTimeOV) / len(solveTimeOV))
        else:
            avrP = avrS = avrSOV = 0
        return avrP, avrS, avrSOV, worstTime, X_star, U_star, goal, x_init, timeout

    def cost_final(self, X_star, U_star):
        return cost_final(X_star, U_star)


def scan_episodes(data_dir="out/data"):
    """montecarlo_cost_ref.py:10-43: {agent: (first_frame, last_frame)} from the ``_cov`` file names."""
    pat = re.compile(r"agent(\d+)_frame(\d+)_cov")
    out = {}
    for name in os.listdir(data_dir):
        m = pat.match(name)
        if not m:
            continue
        a, f = int(m.group(1)), int(m.group(2))
        lo, hi = out.get(a, (f, f))
        out[a] = (min(lo, f), max(hi, f))
    return out


def montecarlo_cost(data_dir="out/data"):
    """montecarlo_cost_ref.py:55-93: averages over the episodes whose last frame did not time out, with
    all episode costs evaluated in ONE batched kernel launch."""
    XL, UU, HH, proc, solve, solveOV, worst = [], [], [], [], [], [], []
    for agent, (lo, hi) in sorted(scan_episodes(data_dir).items()):
        nF = int((hi - lo) / 10 + 1)                                             # :58
        ep = getData_singleEpisode(agent, lo, nF, root=data_dir.rstrip("/") + "/")
        p, s, sov, w, X_star, U_star, _, _, timeout = ep.get_data()
        if timeout[-1] == False:  # noqa: E712  (:66)
            XL.append(np.asarray(X_star[-1], dtype=np.float64).reshape(-1)[:2])
            UU.append(np.asarray([np.asarray(u, dtype=np.float64).reshape(2) for u in U_star]))
            HH.append(len(U_star)); proc.append(p); solve.append(s); solveOV.append(sov); worst.append(w)
    if not XL:
        return dict(n=0)
    dev = torch.device("cuda", torch.cuda.current_device())
    U = np.zeros((len(UU), max(HH), 2))
    for i, u in enumerate(UU):
        U[i, :u.shape[0]] = u
    cost, costU = episode_costs(torch.from_numpy(np.stack(XL)).to(dev), torch.from_numpy(U).to(dev),
                                torch.tensor(HH, dtype=torch.int32, device=dev))
    cost, costU = cost.cpu().numpy(), costU.cpu().numpy()
    n = len(XL)
    return dict(n=n, cost=float(cost.sum() / n), costU=float(costU.sum() / n), costs=cost, costUs=costU,
                processTime=sum(proc) / n, feasibility=n / 50, solveTime=sum(solve) / n,
                solveTimeOV=sum(solveOV) / n, worstTime=sum(worst) / n)
