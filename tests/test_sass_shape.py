"""Code-shape guards on the built library (no GPU needed: cuobjdump reads the sm_100a SASS).

The brute-force kernel's tensor pipe ran at 0.89 of its peak while the MMA issuer rewrote one uniform register pair
with the next descriptor before every tcgen05.mma, and at 0.98 once every descriptor sat in its own pair (DESIGN.md
section 8, tools/ubench/tmem_a_probe.cu).  That property is a matter of what ptxas emits, so it is pinned here: the
eight MMAs of a K loop must be eight consecutive UTCIMMA instructions."""
import functools
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "c_orb_slam_b200", "liborb_gpu.so")


@functools.lru_cache(maxsize=None)
def _dump():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not found")
    if not os.path.exists(LIB):
        pytest.skip("library not built")
    return subprocess.run([exe, "-sass", LIB], capture_output=True, text=True, check=True).stdout


def _sass(kernel_substr):
    txt = _dump()
    for blk in re.split(r"\n\s*Function : ", txt)[1:]:
        name, body = blk.split("\n", 1)
        if kernel_substr in name:
            return [m.group(1).strip() for m in re.finditer(r"/\*[0-9a-f]{4,6}\*/\s+(.*?);", body)]
    pytest.fail("kernel %s not in %s" % (kernel_substr, LIB))


def test_bruteforce_tc5_issues_its_mma_groups_back_to_back():
    ins = _sass("bruteforce_tc5_kernel")
    pos = [k for k, i in enumerate(ins) if "UTCIMMA" in i]
    assert len(pos) == 32, len(pos)  # two stages x two accumulators x 8 K steps
    for g in range(4):
        first, last = pos[8 * g], pos[8 * g + 7]
        between = [i for i in ins[first:last + 1] if "UTCIMMA" not in i]
        # ptxas may schedule a stray instruction into a group; what must not come back is the per-MMA descriptor
        # rewrite (R2UR into the pair the next MMA reads) and the per-MMA election loop (ELECT ... BRA.U.ANY)
        assert len(between) <= 3, "MMA group %d is interleaved with %s" % (g, between)
        assert not any(op in i for i in between for op in ("R2UR", "BRA", "ELECT")), between
        b_desc = [re.findall(r"gdesc\[(UR\d+)\]", ins[k])[1] for k in pos[8 * g:8 * g + 8]]
        assert len(set(b_desc)) == 8, b_desc  # every B descriptor of a group in its own uniform register pair


def test_tensor_core_kernels_use_tcgen05_and_tma():
    assert any("UTCIMMA" in i for i in _sass("blur_tc_kernel"))
    assert any(i.startswith("UTMALDG") or " UTMALDG" in i for i in _sass("blur_tc_kernel"))
    assert any("UTMASTG" in i for i in _sass("blur_tc_kernel"))


def test_staged_kernels_load_their_windows_by_tma():
    """resize_level4_kernel<true> (source bands) and orient_desc_kernel<true> (keypoint windows) stage by TMA; their
    <false> forms, kept for sources TMA cannot address, must not."""
    assert any("UTMALDG" in i for i in _sass("resize_level4_kernelILb1"))
    assert not any("UTMALDG" in i for i in _sass("resize_level4_kernelILb0"))
    assert any("UTMALDG" in i for i in _sass("orient_desc_kernelILb1"))
    assert not any("UTMALDG" in i for i in _sass("orient_desc_kernelILb0"))
