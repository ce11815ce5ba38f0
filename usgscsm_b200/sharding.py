"""Row-band sharding of whole-image batches across the GPUs of one box.

Points are independent and the packed model state is a few KB to MB, so the hot path
shards with NO exchange step: every rank (one process per GPU, torch.distributed) builds
the model on its own device from the same state string and processes a contiguous band of
image rows -- rows map monotonically to time (getImageTime), so a band touches a
contiguous window of the ephemeris tables.  Unordered point lists are split into equal
contiguous chunks.  The only collective is the optional final gather of the outputs.
"""
import torch
import torch.distributed as dist


def row_bands(n_lines, world_size):
    """[(first_line, n_lines_in_band)] for every rank: contiguous, covering, sizes differ by
    at most one line (the first n_lines % world_size ranks get the extra line)."""
    if world_size < 1 or n_lines < 0:
        raise ValueError("bad world size / line count")
    base, extra = divmod(int(n_lines), int(world_size))
    bands, start = [], 0
    for r in range(world_size):
        cnt = base + (1 if r < extra else 0)
        bands.append((start, cnt))
        start += cnt
    return bands


def band_for_rank(n_lines, rank, world_size):
    return row_bands(n_lines, world_size)[rank]


def weighted_row_bands(n_lines, weights):
    """Contiguous row bands sized in proportion to `weights` (one per rank, e.g. the rate each
    rank's host link sustained in a warm-up pass): [(first_line, n_lines_in_band)].  For the
    host-pointer path the ranks of one box share the host's PCIe complex unevenly, and equal
    bands leave the job waiting for the rank with the slowest link.  Every band keeps at least
    one line while there are lines to give; non-positive or non-finite weights count as the mean."""
    world = len(weights)
    good = [w for w in weights if w == w and w > 0 and w != float("inf")]
    mean = sum(good) / len(good) if good else 1.0
    w = [x if (x == x and x > 0 and x != float("inf")) else mean for x in weights]
    total = sum(w)
    edges = [0]
    acc = 0.0
    for r in range(world):
        acc += w[r]
        edges.append(int(round(n_lines * acc / total)))
    edges[-1] = n_lines
    # at least one line per band where possible, edges monotone
    for r in range(1, world + 1):
        edges[r] = max(edges[r], min(edges[r - 1] + 1, n_lines))
    for r in range(world - 1, 0, -1):
        edges[r] = min(edges[r], max(edges[r + 1] - 1, 0)) if n_lines >= world else edges[r]
    return [(edges[r], edges[r + 1] - edges[r]) for r in range(world)]


def chunk_for_rank(n_points, rank, world_size):
    """(offset, count) of this rank's contiguous chunk of an n_points list."""
    return row_bands(n_points, world_size)[rank]


def images_for_rank(n_images, rank, world_size):
    """Whole-image distribution for many small images (config C5): image i goes to rank
    i % world_size."""
    return list(range(rank, n_images, world_size))


def gather_bands(local, n_lines, row_len, group=None):
    """All-gather of row-band shards into the full (n_lines * row_len) array on every rank.
    `local` holds this rank's band, row-major.  Bands may differ by one line, so shards are
    padded to the longest band for the collective (NCCL on GPUs, gloo on CPU)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local.reshape(-1).clone()
    bands = row_bands(n_lines, world)
    longest = max(c for _, c in bands) * row_len
    rank = dist.get_rank(group)
    if local.numel() != bands[rank][1] * row_len:
        raise ValueError("local shard does not match this rank's band")
    padded = torch.zeros(longest, dtype=local.dtype, device=local.device)
    padded[:local.numel()] = local.reshape(-1)
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    return torch.cat([p[:c * row_len] for p, (_, c) in zip(parts, bands)])


class ShardedImage(object):
    """One image, one band per rank.  Construct on every rank with the same model state."""

    def __init__(self, model_state, local_device=None, group=None):
        import usgscsm_b200 as ub
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if local_device is None:
            local_device = self.rank % max(torch.cuda.device_count(), 1)
        self.model = ub.UsgsAstroPlugin(device=local_device).constructModelFromState(model_state)
        size = self.model.getImageSize()
        self.n_lines, self.n_samples = int(size.line), int(size.samp)
        self.first_line, self.band_lines = band_for_rank(self.n_lines, self.rank, self.world)

    def image_to_ground_band(self, height, desiredPrecision=0.001, line_offset=0.5, samp_offset=0.5):
        """Dense-grid imageToGround of this rank's band (pixel centres); `height`: CUDA tensor
        with 1 or band_lines * n_samples entries.  Outputs stay sharded on this rank's GPU."""
        return self.model.imageToGroundGrid(self.first_line + line_offset, samp_offset, 1.0, 1.0,
                                            self.band_lines, self.n_samples, height,
                                            desiredPrecision)

    def ground_to_image_chunk(self, x, y, z, desiredPrecision=0.001):
        """groundToImage of this rank's chunk of ground points (CUDA tensors)."""
        return self.model.groundToImageBatch(x, y, z, desiredPrecision)

    def gather(self, local):
        return gather_bands(local, self.n_lines, self.n_samples, self.group)


def bind_to_gpu_numa_node(device_index):
    """Pin this process (and so its page-locked staging buffers, first-touch) to the CPUs of
    the NUMA node the GPU hangs off.  With one process per GPU the host<->device copies of the
    host-pointer entry points otherwise cross the socket interconnect for half the ranks.
    Returns {"node": n or None, "cpus": count bound to, "why": text}: when the topology gives
    nothing to bind to (a single-node VM reports numa_node = -1) the reason is stated, not hidden."""
    import os
    info = {"node": None, "cpus": None, "why": None}
    try:
        props = torch.cuda.get_device_properties(device_index)
        bus = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        path = "/sys/bus/pci/devices/%s/numa_node" % bus
        with open(path) as f:
            node = int(f.read().strip())
        if node < 0:
            nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")]
            info["why"] = "%s reports %d (%d NUMA node(s) on this host): nothing to bind to" % (path, node, len(nodes))
            return info
        with open("/sys/devices/system/node/node%d/cpulist" % node) as f:
            cpus = set()
            for part in f.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if not allowed:
            info["why"] = "no CPU of node %d is in this process's affinity mask" % node
            return info
        os.sched_setaffinity(0, allowed)
        info.update(node=node, cpus=len(allowed), why="bound")
        return info
    except (OSError, ValueError, AttributeError) as e:
        info["why"] = "topology unreadable: %s" % e
        return info
