"""GPU (-m gpu), collected last: gs_potential_energy_noalloc through the C-ABI.

That entry point serves callers who must not allocate GPU memory (their device memory comes from a
pool / caching allocator, they sit in a latency-critical loop where cudaMalloc's device-wide
synchronisation is not acceptable, or they run at the memory limit): its scratch is a fixed
__device__ array inside the library instead of a cudaMalloc'ed buffer.  Here it is driven the way
such a caller would: every device buffer is owned by the test (torch tensors), the library only
gets raw pointers.  It has to agree with the allocating entry point, gs_energy(), on two calls in a
row -- the second one smaller than the first, so leftovers in the shared scratch would show -- and,
directly, with the fp64 oracle's potential energy."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PE_TOL = 1e-5        # relative, the tolerance the gs_energy tests use for the potential


def pack_pairs(bodies, G):
    """Host-side packing into the layout of DESIGN.md section 3: per pair {x0,x1,y0,y1},{z0,z1,mu0,mu1}."""
    n = bodies.shape[0]
    n_pad = -(-n // 512) * 512
    b = np.zeros((n_pad, 7), dtype=np.float64)
    b[:n] = bodies
    mu = (G * b[:, 6]).astype(np.float32)
    out = np.zeros((n_pad // 2, 8), dtype=np.float32)
    out[:, 0], out[:, 1] = b[0::2, 0], b[1::2, 0]
    out[:, 2], out[:, 3] = b[0::2, 1], b[1::2, 1]
    out[:, 4], out[:, 5] = b[0::2, 2], b[1::2, 2]
    out[:, 6], out[:, 7] = mu[0::2], mu[1::2]
    return out, n_pad


def test_noalloc_potential_matches_gs_energy_twice_in_a_row(gs, oracle):
    import torch
    cases = [(gs.ic_reference_uniform(5000, seed=41), gs.G_SI, 0.0),      # SI units, cutoff rule
             (gs.ic_plummer(1500, seed=42), 1.0, 0.05)]                   # smaller, softened: stale scratch would show
    want = []
    for bodies, G, eps in cases:
        with gs.GravSim(bodies.shape[0], G=G, eps=eps) as sim:            # the allocating entry point
            sim.upload(bodies)
            want.append(sim.energy()[1])
    stream = torch.cuda.Stream()
    got = []
    for bodies, G, eps in cases:                                          # two calls in a row, nothing in between
        packed, n_pad = pack_pairs(bodies, G)
        pos = torch.from_numpy(packed).cuda()                             # caller-owned device memory
        out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        gs.potential_energy_noalloc(pos.data_ptr(), n_pad, G, eps, out.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        got.append(float(out.item()))
    for g, w in zip(got, want):
        assert np.isfinite(g) and abs(g / w - 1.0) <= PE_TOL, (got, want)
    for g, (bodies, G, eps) in zip(got, cases):                           # and against the oracle itself
        assert abs(g / oracle.energy(bodies.astype(np.float64), G=G, eps=eps)[1] - 1.0) <= PE_TOL


def test_noalloc_rejects_what_does_not_fit(gs):
    import torch
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    pos = torch.zeros(512 * 4, dtype=torch.float32, device="cuda")
    too_big = gs.NOALLOC_MAX_BODIES + 512
    with pytest.raises(gs.GravSimError) as e:                             # rejected before anything is touched
        gs.potential_energy_noalloc(pos.data_ptr(), too_big, 1.0, 0.0, out.data_ptr())
    assert e.value.code == -8 and "4194304" in str(e.value)               # GS_ETOOBIG, with the limit spelled out
    with pytest.raises(gs.GravSimError) as e:
        gs.potential_energy_noalloc(pos.data_ptr(), 500, 1.0, 0.0, out.data_ptr())   # not a multiple of 512
    assert e.value.code == -1
    host = np.zeros(2048, dtype=np.float32)                               # not memory of the current device
    with pytest.raises(gs.GravSimError) as e:
        gs.potential_energy_noalloc(host.ctypes.data, 512, 1.0, 0.0, out.data_ptr())
    assert e.value.code == -1
