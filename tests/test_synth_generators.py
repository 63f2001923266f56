"""The three generators of the bench's synthetic genomes must produce the same bytes: numpy (synth.family, used by the
tests), C (oracle.synth_genome / synth_fasta_files, used by the reference arm) and torch (synth.family_device, used by
the GPU arm) -- otherwise the two arms of bench.py would not run the same workload."""
import numpy as np
import pytest
import torch

import oracle
from rabbitkssd_b200 import synth


@pytest.mark.parametrize("seed,length", [(20261106, 100003), (20261106 + 101 * 999, 64), (7, 31)])
def test_c_generator_equals_numpy(seed, length):
    fam = synth.family(seed, 10, length, 0.04, 1)
    for m in range(10):
        assert np.array_equal(oracle.synth_genome(seed, m, 10, length), fam[m][0])


def test_torch_generator_equals_numpy_on_cpu():
    seed, length = 20261106 + 101 * 17, 50021
    fam = synth.family(seed, 10, length, 0.04, 1)
    dev = synth.family_device(torch, seed, 10, length, 0.04, device="cpu")
    for m in range(10):
        assert np.array_equal(dev[m].numpy(), fam[m][0])


def test_fasta_files_equal_write_fasta(tmp_path):
    seed0, length = 20261106, 12345
    oracle.synth_fasta_files(str(tmp_path), seed0, 20, 10, 10, length, 0.04, 2)
    fam = synth.family(seed0 + 101 * 2, 10, length, 0.04, 2)
    for m in (0, 3, 9):
        synth.write_fasta(str(tmp_path / "x.fna"), fam[m], name=f"g{20 + m}.fna")
        assert (tmp_path / "x.fna").read_bytes() == (tmp_path / f"g{20 + m}.fna").read_bytes()


@pytest.mark.gpu
def test_torch_generator_equals_numpy_on_gpu():
    seed, length = 20261106 + 101 * 5, 300007
    fam = synth.family(seed, 10, length, 0.04, 1)
    dev = synth.family_device(torch, seed, 10, length, 0.04, device="cuda")
    for m in range(10):
        assert np.array_equal(dev[m].cpu().numpy(), fam[m][0])
