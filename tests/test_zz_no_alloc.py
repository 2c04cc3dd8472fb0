"""d4b_vec_dot_noalloc (include/d4est_b200.h, csrc/d4b_static.cu): the dot product with its
reduction scratch in a static device array, for callers that cannot call cudaMalloc.  No
counterpart in the reference.  Checked through the C ABI against the existing entry point
d4b_vec_dot (bit-identical: same grid and reduction trees) and against numpy (1e-12 relative,
the tolerance of the other parity tests), on two calls in a row that share the scratch.

This file is named to be collected last."""
import ctypes as C

import numpy as np
import pytest

TOL = 1e-12
MAX_VECS = 64   # D4B_DOT_NOALLOC_MAX_VECS


def _device_vectors(seed, n, nvec):
    import torch

    rng = np.random.default_rng(seed)
    hx, hy = rng.standard_normal(n * nvec), rng.standard_normal(n * nvec)
    dev = torch.device("cuda:0")
    return hx, hy, torch.from_numpy(hx).to(dev), torch.from_numpy(hy).to(dev)


@pytest.mark.gpu
def test_noalloc_dot_without_a_mesh():
    """usable before / without d4b_mesh_create: two calls in a row against numpy"""
    import torch

    from disco4est_b200.runtime import vec_dot_noalloc

    torch.cuda.set_device(0)
    vec_dot_noalloc(None, None, 0, 0, None)      # set-up only: launches nothing
    for seed, n, nvec in [(1, 100003, 5), (2, 777, 2)]:
        hx, hy, x, y = _device_vectors(seed, n, nvec)
        out = torch.full((nvec,), float("nan"), dtype=torch.float64, device=x.device)
        vec_dot_noalloc(x, y, n, nvec, out)
        torch.cuda.synchronize()
        ref = (hx.reshape(nvec, n) * hy.reshape(nvec, n)).sum(axis=1)
        scale = np.abs(hx.reshape(nvec, n) * hy.reshape(nvec, n)).sum(axis=1)
        assert np.all(np.abs(out.cpu().numpy() - ref) <= TOL * scale)


@pytest.mark.gpu
@pytest.mark.parametrize("n,nvec", [(1, 1), (4096, 1), (592 * 256 * 3 + 17, 3), (50021, MAX_VECS), (0, 2)])
def test_noalloc_dot_matches_vec_dot_twice(n, nvec):
    """two calls in a row (the second on other data, on another stream) == d4b_vec_dot bit for bit,
    and the calls change nothing in the device's free memory"""
    import torch

    from disco4est_b200 import Mesh
    from disco4est_b200.brick import brick_desc
    from disco4est_b200.runtime import vec_dot_noalloc

    m = Mesh(brick_desc(2, 2))
    stream = torch.cuda.Stream()
    try:
        data = [_device_vectors(10 + k, n, nvec) for k in range(2)]
        outs = [torch.full((nvec,), float("nan"), dtype=torch.float64, device="cuda:0") for _ in range(2)]
        expect = []
        for hx, hy, x, y in data:
            expect.append([m.L.d4b_vec_dot(m.h, x.data_ptr() + 8 * v * n, y.data_ptr() + 8 * v * n, n) for v in range(nvec)])
        # both modules loaded and both streams used before the free-memory reading
        vec_dot_noalloc(data[0][2], data[0][3], n, 1, outs[0])
        vec_dot_noalloc(data[0][2], data[0][3], n, 1, outs[0], stream=stream.cuda_stream)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        vec_dot_noalloc(data[0][2], data[0][3], n, nvec, outs[0])
        torch.cuda.synchronize()                                  # calls share the scratch: serialise
        vec_dot_noalloc(data[1][2], data[1][3], n, nvec, outs[1], stream=stream.cuda_stream)
        stream.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
        for k in range(2):
            got = outs[k].cpu().numpy()
            assert np.array_equal(got, np.array(expect[k])), (k, got, expect[k])
            hx, hy = data[k][0].reshape(nvec, n), data[k][1].reshape(nvec, n)
            assert np.all(np.abs(got - (hx * hy).sum(axis=1)) <= TOL * np.abs(hx * hy).sum(axis=1))
    finally:
        m.close()


@pytest.mark.gpu
def test_noalloc_dot_rejects_bad_sizes():
    import torch

    from disco4est_b200.runtime import load_library, vec_dot_noalloc

    x = torch.ones(8 * (MAX_VECS + 1), dtype=torch.float64, device="cuda:0")
    out = torch.full((MAX_VECS + 1,), 7.0, dtype=torch.float64, device="cuda:0")
    with pytest.raises(ValueError, match="exceeds D4B_DOT_NOALLOC_MAX_VECS = 64"):
        vec_dot_noalloc(x, x, 8, MAX_VECS + 1, out)
    with pytest.raises(ValueError, match="negative size"):
        vec_dot_noalloc(x, x, -1, 1, out)
    with pytest.raises(ValueError, match="NULL pointer"):
        vec_dot_noalloc(x, None, 8, 1, out)
    torch.cuda.synchronize()
    assert torch.all(out == 7.0)                                 # nothing was launched
    # the error buffer is optional, the code is still returned
    L = load_library()
    assert L.d4b_vec_dot_noalloc(x.data_ptr(), x.data_ptr(), 8, MAX_VECS + 1, out.data_ptr(), None, None, 0) == 2
    small = C.create_string_buffer(8)
    assert L.d4b_vec_dot_noalloc(x.data_ptr(), x.data_ptr(), 8, MAX_VECS + 1, out.data_ptr(), None, small, 8) == 2
    assert small.value == b"d4b_vec"
    # and a good call afterwards still works
    vec_dot_noalloc(x, x, 8, MAX_VECS, out)
    torch.cuda.synchronize()
    assert torch.all(out[:MAX_VECS] == 8.0) and out[MAX_VECS] == 7.0
