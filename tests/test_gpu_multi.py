"""Multi-GPU paths on real GPUs (skipped below 2 visible devices): one process per GPU under torchrun, NCCL.
  * data-parallel training: summed gradients, mean loss, bit-identical parameters on every rank (eager and CUDA-graph replay)
  * sharded tiled inference: bands + seam exchange + label broadcast equal the single-GPU label map"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _run(case, precision, env=None, nproc=2):
    if _ngpus() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    e = dict(os.environ)
    e.update(env or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "tests", "multi_gpu_worker.py"), case, precision]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=e)
    print(r.stdout[-3000:])
    assert r.returncode == 0, r.stderr[-4000:]
    assert "OK " + case in r.stdout


@pytest.mark.parametrize("precision", ["f16", "fp32"])
def test_data_parallel_training_two_ranks(precision):
    _run("dp", precision)


def test_data_parallel_training_two_ranks_eager():
    _run("dp", "f16", env={"AB_NO_GRAPH": "1"})


@pytest.mark.parametrize("precision", ["f16", "fp32"])
def test_sharded_inference_two_ranks(precision):
    _run("infer", precision)


def test_sharded_inference_two_ranks_unbatched_tiles_is_exact():
    """every tile alone: the band protocol itself must reproduce the one-GPU label map (up to exact ties at the seams)"""
    _run("infer", "f16", env={"AB_INFER_BATCH": "1"})
