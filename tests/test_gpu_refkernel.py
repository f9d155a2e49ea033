"""The reference's OWN GPU encoder (src/net/utility/front_top_kernel.cu, compiled unmodified to an sm_100a cubin by
oracle/Makefile) run on this GPU with the reference's launch shapes and host steps (tests/ref_front_top_kernel.py), and
the reference's own equality asserts (src/net/utility/front_top_preprocess.py:213-215, 222-223) applied between ITS
output and the Python-twin mode of this library:

    top.shape == top_gt.shape; top[..., 0:25] == top_gt[..., 0:25]; top[..., 26] == top_gt[..., 26]
    front.shape == front_gt.shape; front == front_gt

(the intensity plane 25 is deliberately not asserted by the reference: its kernel takes the LAST point of a cell in
(x, y, z)-sorted order, the Python encoder the first-highest)."""
import numpy as np
import pytest

from mv3d_b200.params import PYTWIN_KITTI
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("frame", [0, 41])
def test_reference_gpu_kernel_equals_pytwin_mode(torch_cuda, frame):
    torch = torch_cuda
    import ref_front_top_kernel as ref
    if not ref.available():
        pytest.skip("oracle/_ref/front_top_kernel.cubin or cuda-python missing")
    from mv3d_b200.pytwin import PyTwinEncoder
    p = PYTWIN_KITTI
    sweep = hdl64_sweep(frame)
    top = ref.lidar_to_top_cuda(sweep, p)
    front = ref.lidar_to_front_cuda(sweep, p)
    pts, offs = pack_sweeps([sweep])
    enc = PyTwinEncoder(0)
    top_gt, front_gt = enc.encode_device(torch.from_numpy(pts).cuda(), offs, p)
    torch.cuda.synchronize()
    top_gt, front_gt = top_gt[0].cpu().numpy(), front_gt[0].cpu().numpy()
    enc.close()
    zn = p.zn
    assert top.shape == top_gt.shape
    assert np.sum(top[..., 0:zn] != top_gt[..., 0:zn]) == 0          # front_top_preprocess.py:214 (0:25)
    assert np.sum(top[..., zn + 1] != top_gt[..., zn + 1]) == 0      # :215 (26)
    assert front.shape == front_gt.shape
    assert np.sum(front != front_gt) == 0                            # :223
    assert np.count_nonzero(top[..., zn + 1]) > 10000 and np.count_nonzero(front) > 10000  # the comparison is not vacuous
