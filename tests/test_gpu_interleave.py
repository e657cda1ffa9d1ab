"""interleave / deinterleave kernels against the numpy restatement of
compressed::image_algo (image_algo.h:34-65, :84-111):  interleaved[i * nch + c] = plane[c][i].
Bit-exact for every element size and channel count, aligned or not, any length (vector kernel + element tail)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DTYPES = [np.uint8, np.uint16, np.float32, np.float64]


def _planes(dtype, nch, n, seed=0):
    rng = np.random.default_rng(seed + 131 * nch + n)
    raw = rng.integers(0, 256, size=(nch, n * np.dtype(dtype).itemsize), dtype=np.uint8)
    return [raw[c].view(dtype).copy() for c in range(nch)]


def _expected(planes):
    return np.stack(planes, axis=-1).reshape(-1)     # the reference's loop, restated


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("nch", [1, 2, 3, 4, 5, 16])
@pytest.mark.parametrize("n", [1, 15, 16, 17, 255, 4096, 100003])
def test_host_interleave_and_back(dtype, nch, n):
    from compressed_image_b200 import image_algo
    planes = _planes(dtype, nch, n)
    buf = np.zeros(n * nch, dtype=dtype)
    image_algo.interleave(buf, planes)
    assert np.array_equal(buf.view(np.uint8), _expected(planes).view(np.uint8))
    outs = [np.zeros(n, dtype=dtype) for _ in range(nch)]
    image_algo.deinterleave(buf, outs)
    for o, p in zip(outs, planes):
        assert np.array_equal(o.view(np.uint8), p.view(np.uint8))
    outs2 = image_algo.deinterleave(buf, nch)
    assert all(np.array_equal(o.view(np.uint8), p.view(np.uint8)) for o, p in zip(outs2, planes))


@pytest.mark.parametrize("ts", [1, 2, 3, 4, 6, 8, 12, 16])
@pytest.mark.parametrize("shift", [0, 1, 4])
def test_device_any_typesize_and_alignment(ts, shift):
    """Element sizes the vector kernel does not cover and pointers off the 16-byte grid go through the element kernel;
    guard bytes around every output stay untouched."""
    import torch
    from compressed_image_b200 import _b2g
    nch, n = 3, 5000
    rng = np.random.default_rng(ts * 7 + shift)
    planes = rng.integers(0, 256, size=(nch, n * ts), dtype=np.uint8)
    exp = planes.reshape(nch, n, ts).transpose(1, 0, 2).reshape(-1)
    G = 64
    d_planes = [torch.zeros(n * ts + 2 * G, dtype=torch.uint8, device="cuda") for _ in range(nch)]
    for c in range(nch):
        d_planes[c][G + shift:G + shift + n * ts] = torch.from_numpy(planes[c]).cuda()
    d_inter = torch.full((n * nch * ts + 2 * G,), 0xA5, dtype=torch.uint8, device="cuda")
    d_back = [torch.full((n * ts + 2 * G,), 0x5A, dtype=torch.uint8, device="cuda") for _ in range(nch)]
    torch.cuda.synchronize()
    with _b2g.Context() as ctx:
        ctx.interleave_device(ts, [p.data_ptr() + G + shift for p in d_planes], d_inter.data_ptr() + G + shift, n)
        torch.cuda.synchronize()
        ctx.deinterleave_device(ts, d_inter.data_ptr() + G + shift, [p.data_ptr() + G + shift for p in d_back], n)
        torch.cuda.synchronize()
    got = d_inter.cpu().numpy()
    assert np.array_equal(got[G + shift:G + shift + exp.size], exp)
    assert (got[:G + shift] == 0xA5).all() and (got[G + shift + exp.size:] == 0xA5).all()
    for c in range(nch):
        b = d_back[c].cpu().numpy()
        assert np.array_equal(b[G + shift:G + shift + n * ts], planes[c])
        assert (b[:G + shift] == 0x5A).all() and (b[G + shift + n * ts:] == 0x5A).all()


def test_many_channels_cross_launch_boundary():
    """More channels than one launch of the element kernel takes (64): a 70-channel compositing stack."""
    from compressed_image_b200 import image_algo
    planes = _planes(np.uint16, 70, 777)
    buf = np.zeros(70 * 777, dtype=np.uint16)
    image_algo.interleave(buf, planes)
    assert np.array_equal(buf, _expected(planes))
    outs = image_algo.deinterleave(buf, 70)
    assert all(np.array_equal(o, p) for o, p in zip(outs, planes))


def test_full_size_round_trip_and_launch_count():
    """cfg 2 image (8192 x 4096 RGBA half, 256 MiB) on the device: deinterleave -> interleave is the identity, one
    vector-kernel launch each way, and a strided spot check against numpy."""
    import torch
    from compressed_image_b200 import _b2g
    n = 8192 * 4096
    g = torch.Generator(device="cuda").manual_seed(5)
    inter = torch.randint(0, 65536, (n * 4,), dtype=torch.int32, device="cuda", generator=g).to(torch.uint16)
    planes = [torch.zeros(n, dtype=torch.uint16, device="cuda") for _ in range(4)]
    back = torch.zeros_like(inter)
    torch.cuda.synchronize()
    with _b2g.Context() as ctx:
        before = ctx.launches
        ctx.deinterleave_device(2, inter.data_ptr(), [p.data_ptr() for p in planes], n)
        ctx.interleave_device(2, [p.data_ptr() for p in planes], back.data_ptr(), n)   # same stream: ordered
        torch.cuda.synchronize()
        assert ctx.launches - before == 2
    assert torch.equal(back.view(torch.int16), inter.view(torch.int16))
    v = inter.view(torch.int16).view(n, 4)
    for c in range(4):
        assert torch.equal(planes[c].view(torch.int16)[::4099], v[::4099, c])


def test_argument_errors_like_the_reference():
    from compressed_image_b200 import _b2g, image_algo
    a = np.zeros(8, np.uint8)
    with pytest.raises(ValueError, match="No spans"):
        image_algo.interleave(np.zeros(0, np.uint8), [])
    with pytest.raises(ValueError, match="same size"):
        image_algo.interleave(np.zeros(15, np.uint8), [a, a[:7]])
    with pytest.raises(ValueError, match="not large enough"):
        image_algo.interleave(np.zeros(15, np.uint8), [a, a])
    with pytest.raises(ValueError, match="same size"):
        image_algo.deinterleave(np.zeros(16, np.uint8), [a.copy(), a[:7].copy()])
    with pytest.raises(ValueError, match="does not match"):
        image_algo.deinterleave(np.zeros(17, np.uint8), [a.copy(), a.copy()])
    with _b2g.Context() as ctx:
        with pytest.raises(_b2g.B2GError):
            ctx.interleave_device(1, [], 0, 8)
