"""CPU-only checks of the product library: it loads, exports every symbol include/mpccbf.h
declares, validates configs like the reference, builds the same constant tables as the oracle,
and refuses to compute without a device (no fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, swarm


def test_library_exports_every_declared_symbol(M):
    L = M.lib()
    hdr = open(os.path.join(ROOT, "include", "mpccbf.h")).read()
    declared = set(re.findall(r"\b(mpccbf_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(M.EXPORTED_SYMBOLS), declared ^ set(M.EXPORTED_SYMBOLS)
    for s in declared:
        assert hasattr(L, s), s
    assert b"sm_100a" in L.mpccbf_version()


def test_config_struct_layout_matches(M, O):
    assert C.sizeof(M.Config) == C.sizeof(O.Config)
    a, b = M.default_config(), O.base_config()
    for name, _ in M.Config._fields_:
        va, vb = getattr(a, name), getattr(b, name)
        if hasattr(va, "__len__"):
            assert list(va) == list(vb), name
        elif name not in ("max_scenarios", "device", "feas_tol"):
            assert va == vb, name


def test_config_validation_rules(M):
    # common/include/common/parsing.hpp:50-75,121-132,195-213
    assert M.validate(M.default_config())[0] == M.OK
    bad = [dict(num_pieces=0), dict(num_ctrl_pts=0), dict(piece_max_param=0.0), dict(h=0.0), dict(Ts=-1.0),
           dict(k_hor=0), dict(k_hor=20),  # (K-1)h = 1.9 > 3 * 0.5
           dict(v_min=[3, 0, 0]), dict(a_min=[6, 0, 0])]
    for kw in bad:
        rc, msg = M.validate(M.default_config(**kw))
        assert rc == M.ERR_INVALID and msg, kw
    assert M.validate(M.default_config(k_hor=20, num_pieces=4))[0] == M.OK


def test_config_from_reference_json_fixture(M):
    path = os.path.join(ROOT, "tests", "golden", "circle2_config.json")
    cfg, so, sf = M.config_from_json(path)
    assert (cfg.k_hor, cfg.num_pieces, cfg.num_ctrl_pts, cfg.num_agents) == (16, 3, 4, 2)
    assert cfg.d_min == 2.0 and cfg.cbf_horizon == 2 and cfg.impc_iter == 2
    assert np.allclose(so, [[2, 2, 0], [-2, -2, 0]]) and np.allclose(sf, [[-2, -2, 0], [2, 2, 0]])


@pytest.mark.parametrize("K,p", [(16, 3), (20, 4), (10, 2), (40, 8)])
def test_tables_match_oracle(M, O, K, p):
    cfg = M.default_config(k_hor=K, num_pieces=p, device=-1)
    h = M.Handle(cfg)
    oc = O.base_config(k_hor=K, num_pieces=p)
    for name in ("U_basis", "V_basis", "h_samples"):
        assert np.allclose(h.table(name), O.get_table(oc, name), rtol=1e-13, atol=1e-13), name
    st, tg = swarm(1, 3, seed=K)
    qp = O.build_qp(oc, st[0], 1, np.tile(tg[0, 1], K))
    ne = qp["n_eq"]
    scale = np.abs(qp["P"]).max()
    assert np.allclose(h.table("P"), qp["P"], atol=1e-12 * scale, rtol=1e-11)
    assert np.allclose(h.table("E"), qp["A"][:ne], atol=1e-10, rtol=1e-12)
    assert np.allclose(h.table("A_fix"), qp["A"][ne + qp["n_cbf"]:], atol=1e-10, rtol=1e-12)
    q = h.linear_cost(st[0, 1], np.tile(tg[0, 1], K))
    assert np.allclose(q, qp["q"], atol=1e-10 * np.abs(qp["q"]).max())
    # x = Y s + Z w satisfies the equality rows for any w; reduced Hessian is PD (unique minimiser)
    E, Z, Y = h.table("E"), h.table("Z"), h.table("Y")
    assert Z.shape == (h.n, 6)
    assert np.abs(E @ Z).max() < 1e-9
    s = st[0, 1]
    assert np.allclose(E @ (Y @ s), np.concatenate([s, np.zeros(ne - 6)]), atol=1e-9)
    assert np.linalg.eigvalsh(h.table("H_red")).min() > 1.0
    h.close()


def test_no_cpu_fallback(M):
    h = M.Handle(M.default_config(device=-1))
    st, tg = swarm(1, 2)
    with pytest.raises(M.MpcCbfError) as e:
        h.step_batch_host(st, tg)
    assert e.value.code == M.ERR_CUDA and "no CPU fallback" in str(e.value)
    with pytest.raises(M.MpcCbfError):
        h.optimize_one(st[0], 0, np.zeros(48))
    with pytest.raises(M.MpcCbfError):
        h.measure_fp64_tflops()
    h.close()
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(M.MpcCbfError) as e:
            M.Handle(M.default_config(device=0))
        assert e.value.code == M.ERR_CUDA


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "mpc-cbf_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".cpp", ".h", ".hpp", ".py", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle/" not in txt and "pyoracle" not in txt and "liboracle" not in txt, f
