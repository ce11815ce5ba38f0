"""Host-side multi-GPU logic on CPU: the row-band partitioner and the ragged band gather with
world_size 2 and 3 over gloo (the N > 1 path has no data-path collective; the gather is the
only one)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_row_bands_cover_and_balance():
    from usgscsm_b200.sharding import chunk_for_rank, images_for_rank, row_bands
    for n in (0, 1, 7, 52224, 52225, 417792):
        for w in (1, 2, 3, 4, 8):
            b = row_bands(n, w)
            assert len(b) == w and b[0][0] == 0
            assert sum(c for _, c in b) == n
            assert all(b[i][0] + b[i][1] == b[i + 1][0] for i in range(w - 1))
            assert max(c for _, c in b) - min(c for _, c in b) <= 1
    assert chunk_for_rank(10, 1, 4) == (3, 3)
    assert images_for_rank(64, 3, 8) == list(range(3, 64, 8))
    with pytest.raises(ValueError):
        row_bands(10, 0)


def test_weighted_row_bands_cover_and_follow_the_weights():
    """Bands sized to per-rank rates (the host-pointer path on a box whose ranks share the PCIe
    complex unevenly): contiguous, covering, proportional, one line at least where possible."""
    from usgscsm_b200.sharding import row_bands, weighted_row_bands
    rng = np.random.default_rng(0)
    for n in (0, 1, 7, 5064, 52224):
        for w in (1, 2, 3, 8):
            for weights in ([1.0] * w, list(rng.uniform(0.2, 3.0, w)), [0.0] * w,
                            [float("nan")] + [2.0] * (w - 1), [1e-9] + [1.0] * (w - 1)):
                b = weighted_row_bands(n, weights)
                assert len(b) == w and b[0][0] == 0
                assert sum(c for _, c in b) == n and all(c >= 0 for _, c in b)
                assert all(b[i][0] + b[i][1] == b[i + 1][0] for i in range(w - 1))
                if n >= w:
                    assert min(c for _, c in b) >= 1
    assert weighted_row_bands(5064, [1.0] * 8) == row_bands(5064, 8) or \
        max(abs(a[1] - b[1]) for a, b in zip(weighted_row_bands(5064, [1.0] * 8), row_bands(5064, 8))) <= 1
    b = weighted_row_bands(6000, [1.0, 2.0, 3.0])
    assert [c for _, c in b] == [1000, 2000, 3000]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_lines, row_len, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from usgscsm_b200.sharding import band_for_rank, gather_bands
    first, cnt = band_for_rank(n_lines, rank, world)
    rows = torch.arange(first, first + cnt, dtype=torch.float64)[:, None]
    local = (rows * 1000.0 + torch.arange(row_len, dtype=torch.float64)[None, :]).reshape(-1)
    full = gather_bands(local, n_lines, row_len)
    np.save(os.path.join(out_dir, "full_%d.npy" % rank), full.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n_lines", [(2, 9), (2, 10), (3, 10)])
def test_gather_bands_gloo(tmp_path, world, n_lines):
    row_len = 5
    mp.spawn(_worker, args=(world, _free_port(), n_lines, row_len, str(tmp_path)), nprocs=world, join=True)
    want = (np.arange(n_lines)[:, None] * 1000.0 + np.arange(row_len)[None, :]).reshape(-1)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / ("full_%d.npy" % r)), want)
