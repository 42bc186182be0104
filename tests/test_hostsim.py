"""CPU leg: the product's device header (csrc/scn_device.cuh) compiled for the host, checked against the golden
vectors.  This exercises the kernels' control flow (march, refinement quirks, conditioning, NEE) without a GPU; the
GPU leg (test_gpu_parity.py) repeats the same checks through the C ABI on the real kernels."""
import pytest

import golden_util as G
import support as S


@pytest.mark.parametrize("name", sorted(S.MEDIA))
def test_device_header_on_host_vs_golden(name):
    impl = S.CpuImpl("sim", S.MEDIA[name])
    G.check_against_golden(impl, name, G.load(name), exact=False, allow_mismatch=0)
    impl.close()
