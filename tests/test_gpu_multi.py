"""Multi-GPU SyncBatchNorm exchange over NVLink peer memory (csrc/peer.cu; reference
torch.nn.SyncBatchNorm under DDP, Binary_segmentation/train_segmentation.py:199-203).  Needs >= 2 GPUs: the checks live
in scripts/check_syncbn_n2.py (raw exchange vs dist.all_reduce, peer vs NCCL model step, SyncBN over 2 ranks vs one
BatchNorm model over the whole batch) and run here under torchrun."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_syncbn_peer_exchange_two_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29541", os.path.join(ROOT, "scripts", "check_syncbn_n2.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "CHECK_SYNCBN_N2 OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_train_epoch_two_ranks():
    """train_epoch with SyncBatchNorm + GradientAllReduce + FusedAdamW(grad_scale = 1/world): same epoch loss on every
    rank, replicas bit-identical after 12 steps, loss falling (scripts/check_train_n2.py)."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29542", os.path.join(ROOT, "scripts", "check_train_n2.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "CHECK_TRAIN_N2 OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_mae_pretrain_two_ranks():
    """MAE pretraining step under data parallelism (scripts/check_mae_n2.py): bucketed all-reduce of the gradient arena,
    replicas bit-identical after 5 steps, and equal to one process stepping on the concatenated batch."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29543", os.path.join(ROOT, "scripts", "check_mae_n2.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "CHECK_MAE_N2 OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
