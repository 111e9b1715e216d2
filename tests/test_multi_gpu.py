"""Multi-GPU through the C ABI, no torch.distributed on the data path (needs >= 2 B200s; skipped otherwise).

* two processes, one per GPU, exchange the NCCL id through a file (rtw_comm_unique_id / rtw_comm_create), each renders its
  sample shard with rtw_render_sharded, the root's image must equal the single-GPU render of all samples (identical rays;
  sums up to the FP32 addition order of the reduce) and the other rank gets no image;
* the C++ example examples/multi_gpu_render.cpp (the unmodified reference's cornell_box_scene through the adapter, built by
  `make -C oracle example`) writes the same PPM as the single-GPU python path, to the quantisation of the writer."""
import os
import subprocess
import sys

import numpy as np
import pytest

import raytracingweekend_b200 as rtw
from raytracingweekend_b200 import _lib

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, time
import numpy as np
sys.path.insert(0, sys.argv[5])
import raytracingweekend_b200 as rtw
from raytracingweekend_b200.dist import Communicator
rank, world, idf, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4]
if rank == 0:
    cid = Communicator.unique_id()
    open(idf + ".tmp", "wb").write(cid)
    os.rename(idf + ".tmp", idf)
else:
    for _ in range(3000):
        if os.path.exists(idf):
            break
        time.sleep(0.01)
    cid = open(idf, "rb").read()
comm = Communicator(cid, rank, world, rank)
sc = rtw.Scene.named("cornell_box_aluminum", 1.0, device=rank)
for nx, ny, spp, tag in ((64, 64, 33, "a"), (37, 21, 5, "b")):   # 33 and 5 samples do not divide by 2: uneven shards
    acc, st = sc.render_sharded(comm, nx, ny, spp, seed=5)
    assert (acc is not None) == (rank == 0)
    np.save(f"{out}_{tag}_rays{rank}.npy", np.array([st["rays"], st["paths"]], dtype=np.int64))
    if rank == 0:
        np.save(f"{out}_{tag}.npy", acc)
comm.close()
'''


def _n_gpus():
    return _lib.load().rtw_device_count()


@pytest.mark.skipif("_n_gpus() < 2")
def test_two_ranks_render_sharded_equals_one_gpu(tmp_path):
    w = tmp_path / "worker.py"
    w.write_text(WORKER)
    idf, out = str(tmp_path / "nccl.id"), str(tmp_path / "img")
    procs = [subprocess.Popen([sys.executable, str(w), str(r), "2", idf, out, ROOT], cwd=ROOT) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    sc = rtw.Scene.named("cornell_box_aluminum", 1.0)
    for nx, ny, spp, tag in ((64, 64, 33, "a"), (37, 21, 5, "b")):
        one, st = sc.render(nx, ny, spp, seed=5)
        two = np.load(f"{out}_{tag}.npy")
        rays = sum(int(np.load(f"{out}_{tag}_rays{r}.npy")[0]) for r in range(2))
        paths = sum(int(np.load(f"{out}_{tag}_rays{r}.npy")[1]) for r in range(2))
        assert rays == st["rays"] and paths == st["paths"] == nx * ny * spp
        assert np.all(two[..., 3] == spp)
        assert np.allclose(two[..., :3], one[..., :3], rtol=1e-5, atol=1e-5)


@pytest.mark.skipif("_n_gpus() < 2")
def test_cpp_example_on_two_gpus_writes_the_single_gpu_image(tmp_path):
    exe = os.path.join(ROOT, "oracle", "_ref", "multi_gpu_render")
    if not os.path.exists(exe):
        pytest.skip("examples/multi_gpu_render.cpp not built (make -C oracle example needs /root/reference)")
    idf, ppm = str(tmp_path / "nccl.id"), str(tmp_path / "two.ppm")
    nx, ny, spp = 48, 48, 16
    procs = [subprocess.Popen([exe, str(r), "2", idf, str(nx), str(ny), str(spp), ppm], cwd=ROOT) for r in range(2)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    body = open(ppm).read().split()
    assert body[:4] == ["P3", str(nx), str(ny), "255"]
    got = np.array(body[4:], dtype=np.int64).reshape(ny, nx, 3)
    sc = rtw.Scene.named("cornell_box_scene", 1.0)  # the scene the example renders (RayTracingWeekend.cpp:199-204)
    acc, _ = sc.render(nx, ny, spp, seed=0)
    want = (np.float32(255.99).astype(np.float64) * rtw.tonemap(acc)[::-1]).astype(np.int64)
    assert np.abs(got - want).max() <= 1  # the reduce's addition order can move a value across a quantisation step
    assert (got != want).mean() < 0.01
