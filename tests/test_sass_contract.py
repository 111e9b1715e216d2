"""What the compiled kernels must look like for the slot rings of the sorted megakernel to be safe (no GPU needed: reads the SASS).

sorted_kernel.cuh `ring_take` / `ring_put`: each lane has to clear / write its ring position in the very loop iteration in which
its own condition holds.  Written in C++ the store ends up behind the loop's reconvergence barrier (`BSYNC` then `STS`), which
couples the lanes of a warp and let a slot be taken twice on nearly full rings (profiles/r2_ab_phased.txt).  The
loops are PTX for that reason; this test pins what ptxas makes of them: in every `render_sorted` kernel, the 16-bit store that
follows a 16-bit ring load is PREDICATED and sits before the loop's backward branch."""
import os
import re
import shutil
import subprocess

import pytest

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "raytracingweekend_b200", "lib", "librtw_b200.so")
CUOBJDUMP = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"


def _kernels():
    out = subprocess.run([CUOBJDUMP, "-res-usage", LIB], capture_output=True, text=True, timeout=300).stdout
    return sorted(set(re.findall(r"Function (_ZN3rtw13render_sortedI\w+):", out)))


@pytest.mark.skipif(not os.path.exists(CUOBJDUMP) or not os.path.exists(LIB), reason="needs cuobjdump and the built library")
def test_ring_loops_store_inside_the_loop_under_a_predicate():
    kernels = _kernels()
    assert len(kernels) >= 8, kernels  # Cornell, media, textures, three tree variants staged/unstaged, all-features plain + phased
    for k in kernels:
        sass = subprocess.run([CUOBJDUMP, "-sass", "-fun", k, LIB], capture_output=True, text=True, timeout=300).stdout
        ins = [m.group(1).strip() for m in re.finditer(r"/\*[0-9a-f]{4,}\*/\s+(.*?);", sass)]
        loops = 0
        for i, op in enumerate(ins):
            if not op.startswith("LDS.U16"):
                continue
            window = ins[i + 1:i + 8]
            stores = [w for w in window if "STS.U16" in w]
            assert stores, (k, ins[i:i + 8])
            assert stores[0].startswith("@"), f"{k}: ring store not predicated (lanes coupled at the loop exit): {ins[i:i + 8]}"
            j = window.index(stores[0])
            tail = window[j + 1:]
            bra = next((n for n, w in enumerate(tail) if "BRA" in w), None)
            bsync = next((n for n, w in enumerate(tail) if w.startswith("BSYNC")), None)
            assert bra is not None and (bsync is None or bra < bsync), f"{k}: store not inside the loop: {ins[i:i + 10]}"
            loops += 1
        assert loops >= 2, (k, loops)  # ring_take and ring_put (inlined once or more)
