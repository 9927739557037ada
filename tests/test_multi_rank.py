"""N > 1 path: world_size-2 gloo run on CPU (host-side partition plumbing) and, on a box with >= 2 GPUs, the real
NCCL path against the single-rank result."""
import os
import socket
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _launch(nproc, args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "mgpu_worker.py")] + args
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    return r.returncode, r.stdout + r.stderr


def test_partition_plumbing_gloo_world2():
    rc, out = _launch(2, ["gloo"])
    assert rc == 0 and "MGPU-OK gloo" in out, out[-3000:]


@pytest.mark.gpu
@pytest.mark.parametrize("nproc", [2, 4, 8])
@pytest.mark.parametrize("args", [["nccl", "stretched"], ["nccl", "stretched", "periodic"], ["nccl", "terrain"]])
def test_n_ranks_match_one_rank_nccl(args, nproc):
    """2 ranks: every rank has ONE neighbour (and, periodic, the same neighbour on both sides); 4 and 8 ranks: interior
    ranks with two distinct peers -- two IPC mappings, independent ready/ack counters per side (halo_p2p.cu)."""
    import torch
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    rc, out = _launch(nproc, args)
    assert rc == 0 and "MGPU-OK nccl" in out, out[-4000:]


@pytest.mark.gpu
@pytest.mark.parametrize("nproc", [2, 4])
def test_advection_predictor_on_y_slab_ranks(nproc):
    """predictScalar (PPM / PLM / CTU, both limiter families) on 2 and 4 ranks, bit-exact against the numpy oracle of the
    whole periodic domain"""
    import torch
    if torch.cuda.device_count() < nproc:
        pytest.skip(f"needs {nproc} GPUs")
    rc, out = _launch(nproc, ["advect"])
    assert rc == 0 and "MGPU-OK nccl" in out, out[-4000:]

