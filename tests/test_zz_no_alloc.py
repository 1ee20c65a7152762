"""rnnt_b200_lattice_static: the lattice entry point for callers that cannot allocate (alpha / beta workspace in a
file-scope device array inside the library); no counterpart in the
reference. Collected LAST on purpose: a failure here hides no other test.

The GPU test calls the new entry point through the C ABI (ctypes) and checks its outputs against the existing
rnnt_b200_lattice on the same planes, on two calls in a row with different shapes (the second call finds the first
one's contents in the shared scratch)."""
import ctypes

import pytest
import torch

from rnnt_transducer_b200 import _lib

PLANE_MAX = 1 << 21   # RNNT_B200_STATIC_PLANE_ELEMS
B_MAX = 1024          # RNNT_B200_STATIC_MAX_B


def test_static_entry_rejects_sizes_beyond_its_fixed_scratch():
    # argument checks run before anything touches a device: no GPU needed
    L = _lib.lib()
    a = ctypes.c_void_p(256)
    args = lambda B, T, U1: (a, a, a, a, B, T, U1, a, a, a, None, None)
    assert L.rnnt_b200_plane_elems(32, 500, 101) <= PLANE_MAX          # the headline shape fits
    assert L.rnnt_b200_lattice_static(*args(64, 500, 101)) == -5        # RNNT_B200_E_WORKSPACE: too many cells
    assert L.rnnt_b200_lattice_static(*args(B_MAX + 1, 1, 1)) == -5     # too many utterances
    assert L.rnnt_b200_lattice_static(*args(1, 4, 2000)) == -2          # U1 > 1024: shape, as rnnt_b200_lattice
    assert L.rnnt_b200_lattice_static(None, a, a, a, 1, 4, 2, a, a, a, None, None) == -1
    assert L.rnnt_b200_status_string(-5) == b"workspace too small"


def _planes(L, B, T, U, V, seed):
    """lp_blank / lp_label planes of random logits through rnnt_b200_logits_to_planes (ragged lengths)."""
    from tests.helpers import synth_labels, synth_lengths
    dev = "cuda:0"
    g = torch.Generator().manual_seed(seed)
    tl, ul = synth_lengths(B, T, U, "edge" if B >= 3 else "ragged", g)
    y = synth_labels(B, U, V, ul, g)
    x = torch.randn(B, T, U + 1, V, generator=g).to(dev)
    y, tl, ul = y.to(dev), tl.to(dev), ul.to(dev)
    n = L.rnnt_b200_plane_elems(B, T, U + 1)
    lse, lpb, lpl = (torch.empty(n, device=dev) for _ in range(3))
    st = torch.cuda.current_stream().cuda_stream
    rc = L.rnnt_b200_logits_to_planes(x.data_ptr(), y.data_ptr(), tl.data_ptr(), ul.data_ptr(), B, T, U + 1, V, 0,
                                      lse.data_ptr(), lpb.data_ptr(), lpl.data_ptr(), st)
    assert rc == 0, rc
    return lpb, lpl, tl, ul, n


@pytest.mark.gpu
def test_static_lattice_matches_the_workspace_entry_point_on_two_calls_in_a_row():
    L = _lib.lib()
    dev = "cuda:0"
    torch.cuda.set_device(0)
    st = torch.cuda.current_stream().cuda_stream
    # second shape smaller than the first: its planes overlap what the first call left in the shared scratch
    for (B, T, U, V, seed) in [(4, 57, 19, 31, 1), (3, 23, 40, 17, 2)]:
        lpb, lpl, tl, ul, n = _planes(L, B, T, U, V, seed)
        U1 = U + 1
        # existing entry point, caller-owned float64 workspace
        alpha = torch.empty(n + B, dtype=torch.float64, device=dev)
        beta = torch.empty(n + B, dtype=torch.float64, device=dev)
        cb0, cl0 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        loss0 = torch.zeros(B, device=dev)
        l64_0 = torch.zeros(B, dtype=torch.float64, device=dev)
        rc = L.rnnt_b200_lattice(lpb.data_ptr(), lpl.data_ptr(), tl.data_ptr(), ul.data_ptr(), B, T, U1,
                                 alpha.data_ptr(), beta.data_ptr(), cb0.data_ptr(), cl0.data_ptr(),
                                 loss0.data_ptr(), l64_0.data_ptr(), st)
        assert rc == 0, rc
        # new entry point: inputs and outputs only (allocated here, before the call; the call itself allocates nothing)
        cb1, cl1 = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        loss1 = torch.zeros(B, device=dev)
        l64_1 = torch.zeros(B, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = L.rnnt_b200_lattice_static(lpb.data_ptr(), lpl.data_ptr(), tl.data_ptr(), ul.data_ptr(), B, T, U1,
                                        cb1.data_ptr(), cl1.data_ptr(), loss1.data_ptr(), l64_1.data_ptr(), st)
        assert rc == 0, (rc, L.rnnt_b200_status_string(rc))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert bool(torch.isfinite(loss1).all()) and float(loss1.min()) > 0
        # the fp32-path tolerances of tests/test_dropin_gpu.py: 1e-5 relative on the loss, 1e-4 absolute on the
        # gradient-side quantities (the same kernels run, so the results are in fact bit-identical)
        assert float(((loss1 - loss0).abs() / loss0.abs()).max()) < 1e-5
        assert float(((l64_1 - l64_0).abs() / l64_0.abs()).max()) < 1e-5
        assert float((cb1 - cb0).abs().max()) < 1e-4 and float((cl1 - cl0).abs().max()) < 1e-4
        assert torch.equal(loss1, loss0) and torch.equal(cb1, cb0) and torch.equal(cl1, cl0)
