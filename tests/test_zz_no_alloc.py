"""crt_postprocess_device_noalloc (include/crt.h): the depth-of-field path with its scratch in a fixed
`static __device__` array of the library instead of cudaMalloc'd memory.  It has no counterpart in the
reference, so the check is against the existing entry
point crt_postprocess_device on the same device buffers -- BIT-EXACT, the tolerance of
tests/test_gpu_post.py: both run the same kernels in the same order -- on two calls in a row that share
the array (the second on a smaller frame with other flags, so it sees the first call's leftovers).

Named test_zz_* so that pytest collects it last: `pytest -x` must not lose the other
tests to a failure here.  Passes alone in a fresh process and after the rest of the suite."""
import numpy as np
import pytest

from c_raytracer_b200 import device, host
from test_oracle_post import seeded_frame

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    if device.device_count() < 1:
        pytest.fail("no CUDA device: the post-processing path has no CPU fallback")
    return 0


def _both(torch, w, h, seed, flags, check_free=False):
    """-> (frame of crt_postprocess_device_noalloc, frame of crt_postprocess_device, the input), from the same input"""
    raster, z = seeded_frame(w, h, seed)
    o = host.post_options(*flags)
    d_z = torch.from_numpy(z).to("cuda:0")
    d_want = torch.from_numpy(raster).to("cuda:0")
    d_got = d_want.clone()
    stream = torch.cuda.Stream()
    torch.cuda.synchronize()
    device.postprocess_device(o, w, h, d_want.data_ptr(), d_z.data_ptr(), 0, stream.cuda_stream)
    stream.synchronize()
    free_before = torch.cuda.mem_get_info(0)[0]
    device.postprocess_device_noalloc(o, w, h, d_got.data_ptr(), d_z.data_ptr(), 0, stream.cuda_stream)
    stream.synchronize()
    free_after = torch.cuda.mem_get_info(0)[0]
    if check_free:
        # nothing was allocated or released on the device by the call (the first call of a process may load the module)
        assert free_after == free_before
    np.testing.assert_array_equal(d_z.cpu().numpy(), z)  # the z-buffer is read only
    return d_got.cpu().numpy(), d_want.cpu().numpy(), raster


def test_two_calls_in_a_row_equal_the_allocating_entry_point(gpu):
    torch = pytest.importorskip("torch")
    got, want, before = _both(torch, 640, 360, 51, ("-b", "1.3", "--dof", "2.2", "-13", "--mist", "3", "5", "quad", "0.2", "0.3", "0.4"))
    np.testing.assert_array_equal(got, want)
    assert not np.array_equal(got, before)
    # second call: smaller frame, ragged width, smaller circles of confusion, on the array the first call left behind
    got, want, before = _both(torch, 333, 187, 52, ("--dof", "0.9", "-5"), check_free=True)
    np.testing.assert_array_equal(got, want)
    assert not np.array_equal(got, before)
    # and a full-HD frame, the largest standard size the array holds
    got, want, _ = _both(torch, 1920, 1080, 53, ("--dof", "0.6", "-3"), check_free=True)
    np.testing.assert_array_equal(got, want)


def test_limits_and_errors(gpu):
    torch = pytest.importorskip("torch")
    tiny = torch.zeros(16, dtype=torch.float32, device="cuda:0")
    o = host.post_options("--dof", "0.9", "-5")
    # beyond the maximum: refused before anything touches the pointers
    with pytest.raises(device.CrtError, match="CRT_POST_NOALLOC_MAX_PIXELS"):
        device.postprocess_device_noalloc(o, 2048, 1025, tiny.data_ptr(), tiny.data_ptr())
    with pytest.raises(device.CrtError, match="CRT_POST_NOALLOC_MAX_PIXELS"):
        device.postprocess_device_noalloc(o, 3840, 2160, tiny.data_ptr(), tiny.data_ptr())
    # within the pixel limit but too many rows / segments for the array (one-pixel-wide frame)
    with pytest.raises(device.CrtError, match="CRT_POST_NOALLOC_SCRATCH_BYTES"):
        device.postprocess_device_noalloc(o, 1, device.NOALLOC_MAX_PIXELS, tiny.data_ptr(), tiny.data_ptr())
    # not the depth-of-field path
    with pytest.raises(device.CrtError, match="CRT_POST_DOF"):
        device.postprocess_device_noalloc(host.post_options("-b", 2), 2, 2, tiny.data_ptr(), tiny.data_ptr())
    with pytest.raises(device.CrtError, match="NULL"):
        device.postprocess_device_noalloc(o, 2, 2, 0, tiny.data_ptr())
    with pytest.raises(device.CrtError):
        device.postprocess_device_noalloc(o, 2, 2, tiny.data_ptr(), tiny.data_ptr(), device=99)
    # the library still works after the refusals
    torch.cuda.synchronize()
    got, want, _ = _both(torch, 31, 57, 54, ("--dof", "1.5", "-8"))
    np.testing.assert_array_equal(got, want)
