"""The oracle (oracle/hydragnn_ref.py) against golden vectors produced by the REAL
reference modules (tests/golden/make_golden.py).  CPU only."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
import make_golden as mg  # noqa: E402  (only its pure helpers are used; nothing reads /root/reference)

from oracle import hydragnn_ref as ora  # noqa: E402

ORACLE_CLASSES = {"SCFStack": ora.SCFStack, "EGCLStack": ora.EGCLStack,
                  "PAINNStack": ora.PAINNStack, "PNAStack": ora.PNAStack}
TOL = dict(rtol=1e-5, atol=1e-6)


def test_small_functions(golden):
    f = golden["functions"]
    v, l = ora.get_edge_vectors_and_lengths(f["pos"], f["edge_index"], f["shifts"])
    assert torch.equal(v, f["vec"]) and torch.equal(l, f["len"])
    v, l = ora.get_edge_vectors_and_lengths(f["pos"], f["edge_index"], f["shifts"], normalize=True, eps=1.0)
    assert torch.equal(v, f["vec_norm"]) and torch.equal(l, f["len_norm"])
    assert torch.equal(ora.unsorted_segment_mean(f["data"], f["edge_index"][0], 7), f["seg_mean"])
    assert torch.equal(ora.unsorted_segment_sum(f["data"], f["edge_index"][0], 7), f["seg_sum"])
    assert torch.equal(ora.sinc_expansion(f["dist"], 5, 3.0), f["sinc"])
    assert torch.equal(ora.cosine_cutoff(f["dist"], 3.0), f["cutoff"])


import make_golden_train as mgt  # noqa: E402


@pytest.mark.parametrize("name", sorted(mg.MODEL_CASES) + sorted(mgt.TRAIN_CASES))
def test_models_match_reference(golden, name):
    fx = golden[name]
    kind, cfg = fx["kind"], fx["cfg"]
    model = mg.build(ORACLE_CLASSES, kind, cfg)
    model.load_state_dict(fx["state"], strict=True)   # same parameter names as the reference
    if cfg.get("eval"):
        model.eval()
    out = mg.run_model(model, fx["batch"], cfg)
    ref = fx["out"]
    torch.testing.assert_close(out["loss"], ref["loss"], **TOL)
    k = 0
    while f"pred{k}" in ref:
        torch.testing.assert_close(out[f"pred{k}"], ref[f"pred{k}"], **TOL)
        k += 1
    assert set(out["grads"]) == set(ref["grads"])
    for n, g in ref["grads"].items():
        torch.testing.assert_close(out["grads"][n], g, rtol=1e-4, atol=1e-6)
    if cfg.get("forces"):
        torch.testing.assert_close(out["pos_grad"], ref["pos_grad"], rtol=1e-4, atol=1e-6)
    for n, b in ref.get("buffers", {}).items():     # train mode: BatchNorm running statistics after the step
        torch.testing.assert_close(dict(model.named_buffers())[n], b, **TOL)


@pytest.mark.parametrize("case", ["h2_loop0", "h2_loop1", "cr_loop0", "cr_loop1", "tri_loop0",
                                  "tri_loop1", "tri_cap_loop0", "tri_cap_loop1", "h2_big_loop0"])
def test_radius_graph_pbc_matches_reference(golden, case):
    c = golden["pbc"][case]
    ei, S, d = ora.radius_graph_pbc(c["pos"].numpy(), c["cell"].numpy(), c["pbc"].numpy(), c["r"],
                                    c["cap"], c["loop"])
    shifts = torch.from_numpy(S).to(torch.float32) @ c["cell"]
    ref_ei, ref_sh = c["edge_index"], c["edge_shifts"]
    assert ei.shape[1] == ref_ei.shape[1]
    # canonical sort: (dst, src, shift vector) -- lengths tie on perfect lattices
    def canon(e, s):
        e = np.asarray(e); s = np.round(np.asarray(s), 4)
        order = np.lexsort((s[:, 2], s[:, 1], s[:, 0], e[0], e[1]))
        return e[:, order], s[order]
    a_e, a_s = canon(ei, shifts.numpy())
    b_e, b_s = canon(ref_ei.numpy(), ref_sh.numpy())
    if "cap" in case:   # capped sets are only defined up to distance ties; compare multiset of dst
        assert np.array_equal(np.sort(a_e[1]), np.sort(b_e[1]))
    else:
        assert np.array_equal(a_e, b_e)
        assert np.allclose(a_s, b_s, atol=1e-4)


def test_neighbour_count_known_answers(golden):
    """tests/test_periodic_boundary_conditions.py:82-127 of the reference."""
    for case, per_atom in [("h2_loop0", 1), ("h2_loop1", 2), ("cr_loop0", 14), ("cr_loop1", 15)]:
        c = golden["pbc"][case]
        ei, S, d = ora.radius_graph_pbc(c["pos"].numpy(), c["cell"].numpy(), c["pbc"].numpy(), c["r"],
                                        c["cap"], c["loop"])
        assert ei.shape[1] == per_atom * c["pos"].shape[0]
        assert (d >= 0).all() and (d <= c["r"]).all()
