"""GPU: the device-resident episode driver (mpccbf_run_episode) against (i) the host-driven loop over
the same C-ABI calls, (ii) the CPU oracle stepped in closed loop, (iii) a numpy restatement of the
reference's instance_success metric (experiments/python/metrics/collision_check.py:48-79)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, swarm

pytestmark = pytest.mark.gpu


def instance_success(traj, goals, radius, coll_radius):
    """collision_check.py:48-79, circle shape. traj [N][T][6]. Returns (success, finish) with
    finish = -1 for float('inf')."""
    N, T = traj.shape[:2]
    reach = [False] * N
    for t in range(T):
        if all(reach):
            return True, max(0, t - 1)
        for i in range(N):
            if np.linalg.norm(traj[i, t, :2] - goals[i][:2]) <= radius:
                reach[i] = True
            for j in range(i + 1, N):
                if np.hypot(*(traj[j, t, :2] - traj[i, t, :2])) <= 2 * coll_radius:
                    return False, -1
    return True, T


def run(M, cfg, st, tg, steps, goal_r, coll_r, **kw):
    import torch
    S, N = st.shape[:2]
    h = M.Handle(cfg)
    nsub = int(cfg.h / cfg.Ts)
    d_st, d_tg = torch.tensor(st, device="cuda"), torch.tensor(tg, device="cuda")
    out = dict(success=torch.zeros(S, dtype=torch.int32, device="cuda"), finish=torch.zeros(S, dtype=torch.int32, device="cuda"),
               min_dist=torch.zeros(S, dtype=torch.float64, device="cuda"), failed=torch.zeros(S, dtype=torch.int32, device="cuda"),
               traj=torch.zeros((S, N, steps * nsub, 6), dtype=torch.float64, device="cuda"))
    h.run_episode(S, d_st, d_tg, steps, goal_r, coll_r, out["success"], out["finish"], out["min_dist"], out["failed"], out["traj"],
                  **kw)
    torch.cuda.synchronize()
    res = {k: v.cpu().numpy() for k, v in out.items()}
    res["final"] = d_st.cpu().numpy()
    return h, res


def test_episode_matches_oracle_closed_loop(M, O):
    # 2 scenarios x 4 agents, 12 control steps: every sub-step state against the oracle stepped in
    # closed loop with the reference driver's rules (MPCCBFFormationControl_example.cpp:140-231)
    S, N, steps = 2, 4, 12
    cfg = M.default_config(k_hor=16, num_pieces=3, d_min=0.8, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=5)
    st[:, :, :2] *= 0.5
    tg[:, :, :2] *= 0.5
    h, res = run(M, cfg, st, tg, steps, 1.0, 0.2)
    oc = O.base_config(k_hor=16, num_pieces=3, d_min=0.8, num_agents=N)
    s = st.copy()
    prev = [[None] * N for _ in range(S)]
    et = np.zeros((S, N))
    maxp = 3 * 0.5
    for step in range(steps):
        r = O.step_batch(oc, s, tg, threads=2, prec=1)
        for sc in range(S):
            for a in range(N):
                if r["n_curves"][sc, a] > 0:
                    prev[sc][a], et[sc, a] = r["ctrl_pts"][sc, a].copy(), 0.0
        for ti in range(1, 11):
            for sc in range(S):
                for a in range(N):
                    if prev[sc][a] is None:
                        s[sc, a, 3:] = 0.0
                        continue
                    t = min(et[sc, a] + 0.01 * ti, maxp)
                    s[sc, a, :3] = O.eval_curve(oc, prev[sc][a], t, 0)
                    s[sc, a, 3:] = O.eval_curve(oc, prev[sc][a], t, 1)
            assert np.abs(res["traj"][:, :, 10 * step + ti - 1, :] - s).max() < 1e-6, (step, ti)
        et = np.minimum(et + 0.1, maxp)
    assert np.abs(res["final"] - s).max() < 1e-6
    h.close()


def test_episode_metrics_and_host_loop(M):
    import torch
    # 48 ring-swap scenarios of 8 agents, 70 control steps: metric against the numpy restatement on the
    # dumped trajectory, final states against the host-driven loop over step_batch + curve following
    S, N, steps = 48, 8, 70
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=11)
    st[:, :, :2] *= 0.6
    tg[:, :, :2] *= 0.6
    tg[::2, :, :2] = st[::2, :, :2] + [2.5, 1.0]     # every other scenario: a formation translation (no conflicts)
    coll_r = 0.36
    h, res = run(M, cfg, st, tg, steps, 1.0, coll_r)
    nfin, nfail = 0, 0
    for s in range(S):
        ok, fin = instance_success(res["traj"][s], tg[s], 1.0, coll_r)
        assert bool(res["success"][s]) == ok and res["finish"][s] == fin, (s, ok, fin, res["success"][s], res["finish"][s])
        d = np.linalg.norm(res["traj"][s][:, None, :, :2] - res["traj"][s][None, :, :, :2], axis=-1)
        d[np.arange(N), np.arange(N)] = np.inf
        assert abs(d.min() - res["min_dist"][s]) < 1e-12
        nfin += int(ok and fin < steps * 10)
        nfail += int(not ok)
    assert nfin >= S // 2      # the translations complete inside the episode
    assert np.array_equal(res["final"], res["traj"][:, :, -1, :])
    # same episode without the trajectory dump and metrics pointers: identical final states
    d_st = torch.tensor(st, device="cuda")
    h.run_episode(S, d_st, torch.tensor(tg, device="cuda"), steps, 1.0, coll_r)
    torch.cuda.synchronize()
    assert np.array_equal(d_st.cpu().numpy(), res["final"])
    h.close()


def test_episode_two_robot_config(M):
    # BASELINE configs[0]: the shipped 2-robot swap (circle2_config.json), 200 control steps
    cfg, so, sf = M.config_from_json(os.path.join(GOLDEN, "circle2_config.json"), max_scenarios=1)
    assert cfg.num_agents == 2
    st = np.zeros((1, 2, 6))
    st[0, :, :3] = so[:2]
    tg = sf[None, :2, :].copy()
    h, res = run(M, cfg, st, tg, 200, 1.0, 0.3)
    assert np.abs(res["final"][0, :, :2] - tg[0, :, :2]).max() < 0.05
    assert res["traj"].shape == (1, 2, 2000, 6)
    assert res["failed"][0] > 0          # the exact head-on swap has infeasible steps
    h.close()


def test_episode_gauss_seidel_matches_host_driver(M, tmp_path):
    # the reference's loop order: the host driver (formation_example --mode gauss_seidel, one optimize_one call
    # per robot) and the device episode in MPCCBF_ORDER_GAUSS_SEIDEL must produce the same states.json
    import json
    import subprocess
    from conftest import ROOT
    host = os.path.join(ROOT, "mpc-cbf_b200", "host")
    subprocess.run(["make", "-C", host], check=True, stdout=subprocess.DEVNULL)
    out = tmp_path / "gs.json"
    r = subprocess.run([os.path.join(host, "formation_example"), "--config_file", os.path.join(GOLDEN, "circle2_config.json"),
                        "--write_filename", str(out), "--sim_runtime", "20", "--mode", "gauss_seidel"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    js = json.load(open(out))
    ref = np.array([js["robots"][str(i)]["states"] for i in range(2)])[:, :2000]     # [2][2000][6]
    assert ref.shape == (2, 2000, 6)
    cfg, so, sf = M.config_from_json(os.path.join(GOLDEN, "circle2_config.json"), max_scenarios=1)
    st = np.zeros((1, 2, 6))
    st[0, :, :3] = so[:2]
    h, res = run(M, cfg, st, sf[None, :2, :].copy(), 200, 1.0, 0.3, order=M.ORDER_GAUSS_SEIDEL)
    assert np.abs(res["traj"][0] - ref).max() < 1e-6
    h.close()


def test_episode_gauss_seidel_matches_oracle(M, O):
    # a pulled-in 4-agent swarm where safety rows bind: the Gauss-Seidel episode against the oracle stepped
    # robot by robot (robot i optimises on the states robots < i have already advanced to), and it must
    # differ from the Jacobi episode
    S, N, steps = 1, 4, 10
    cfg = M.default_config(k_hor=16, num_pieces=3, d_min=0.8, num_agents=N, max_scenarios=S)
    from parity import dense_scenarios
    st, tg = dense_scenarios(S, N, 3, 0.8, spread=1.2, vmax=1.0)
    h, gs = run(M, cfg, st, tg, steps, 1.0, 0.2, order=M.ORDER_GAUSS_SEIDEL)
    h2, jac = run(M, cfg, st, tg, steps, 1.0, 0.2)
    assert np.abs(gs["traj"] - jac["traj"]).max() > 1e-4
    oc = O.base_config(k_hor=16, num_pieces=3, d_min=0.8, num_agents=N)
    s = st[0].copy()
    prev, et = [None] * N, np.zeros(N)
    for step in range(steps):
        for a in range(N):
            r = O.optimize(oc, s, a, np.tile(tg[0, a], 16), prec=1)
            if r["n_curves"] > 0:
                prev[a], et[a] = r["curves"][-1].copy(), 0.0
            for ti in range(1, 11):
                if prev[a] is None:
                    s[a, 3:] = 0.0
                else:
                    t = min(et[a] + 0.01 * ti, 1.5)
                    s[a, :3] = O.eval_curve(oc, prev[a], t, 0)
                    s[a, 3:] = O.eval_curve(oc, prev[a], t, 1)
                assert np.abs(gs["traj"][0, a, 10 * step + ti - 1] - s[a]).max() < 1e-6, (step, a, ti)
            et[a] = min(et[a] + 0.1, 1.5)
    h.close()
    h2.close()


def test_episode_noise_statistics(M):
    # seeded N(0, std) on every component at every sub-step (math::addRandomNoise): deterministic per seed,
    # right moments on the first control step (where the clean trajectory is the same)
    S, N = 64, 8
    cfg = M.default_config(k_hor=20, num_pieces=4, d_min=0.8, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=2)
    h0, clean = run(M, cfg, st, tg, 1, 1.0, 0.2)
    h1, n1 = run(M, cfg, st, tg, 1, 1.0, 0.2, pos_std=0.01, vel_std=0.05, seed=123)
    h2, n2 = run(M, cfg, st, tg, 1, 1.0, 0.2, pos_std=0.01, vel_std=0.05, seed=123)
    h3, n3 = run(M, cfg, st, tg, 1, 1.0, 0.2, pos_std=0.01, vel_std=0.05, seed=124)
    assert np.array_equal(n1["traj"], n2["traj"]) and not np.array_equal(n1["traj"], n3["traj"])
    d = n1["traj"] - clean["traj"]            # [S][N][10][6]
    for sl, std in ((slice(0, 3), 0.01), (slice(3, 6), 0.05)):
        x = d[..., sl].ravel() / std
        assert abs(x.mean()) < 4 / np.sqrt(x.size) and abs(x.std() - 1.0) < 0.03
        assert abs(np.mean(x ** 4) - 3.0) < 0.3                 # Gaussian kurtosis
    # components and sub-steps are uncorrelated
    c = np.corrcoef(d[..., 0].ravel(), d[..., 1].ravel())[0, 1]
    assert abs(c) < 0.05
    for h in (h0, h1, h2, h3):
        h.close()


def instance_success_box(traj, goals, radius, half):
    """collision_check.py:10-79 with shape_type == "box" (corner = centre - half / 2, width 2 * half)."""
    N, T = traj.shape[:2]
    cx, cy = half
    reach = [False] * N
    for t in range(T):
        if all(reach):
            return True, max(0, t - 1)
        for i in range(N):
            if np.linalg.norm(traj[i, t, :2] - goals[i][:2]) <= radius:
                reach[i] = True
            for j in range(i + 1, N):
                xA, yA = traj[i, t, 0] - cx / 2, traj[i, t, 1] - cy / 2
                xB, yB = traj[j, t, 0] - cx / 2, traj[j, t, 1] - cy / 2
                if xA < xB + 2 * cx and xA + 2 * cx > xB and yA < yB + 2 * cy and yA + 2 * cy > yB:
                    return False, -1
    return True, T


def test_episode_box_collision_shape_of_the_reference_configs(M):
    # base_config.json / circle{2,8}.json carry "aligned_box": [0.2, 0.2, 0] next to "radius": the metric then
    # uses the rectangle test (collision_check.py:93-101). Shape parsed from the golden config; box sizes from
    # the file's own to one that makes some scenarios collide.
    o = M.episode_opts_from_json(os.path.join(GOLDEN, "circle8_config.json"))
    assert o.collision_shape == M.SHAPE_BOX and list(o.box_half) == [0.2, 0.2] and o.goal_radius == 1.0
    S, N, steps = 12, 6, 40
    cfg = M.default_config(k_hor=16, num_pieces=3, d_min=0.8, num_agents=N, max_scenarios=S)
    st, tg = swarm(S, N, seed=9)
    st[:, :, :2] *= 0.45
    tg[:, :, :2] *= 0.45
    st[0, :, :2] *= 3.0 / 0.45       # scenario 0: spread out and already at its goals, so it succeeds for every box
    st[0, :, 3:] = 0.0
    tg[0] = st[0, :, :3]
    seen = set()
    for half in ((0.2, 0.2), (0.45, 0.3), (0.7, 0.7)):
        h, res = run(M, cfg, st, tg, steps, 1.0, 0.2, box_half=half)
        for s in range(S):
            ok, fin = instance_success_box(res["traj"][s], tg[s], 1.0, half)
            assert bool(res["success"][s]) == ok and int(res["finish"][s]) == fin, (half, s)
            seen.add(ok)
        h.close()
    assert seen == {True, False}
