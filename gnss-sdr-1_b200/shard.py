"""Multi-GPU sharding of the correlator path: one process per GPU, no exchange step inside a unit of work.

* Acquisition units are (PRN, Doppler-bin range).  PRNs are dealt round-robin over the ranks; when there are fewer PRNs than
  ranks each PRN's Doppler grid is additionally cut into contiguous bin ranges.  A bin range [lo, hi) of the reference grid
  ``doppler = -doppler_max + step * i`` (pcps_acquisition.cc:353-359) is searched through the unchanged C ABI as a smaller grid
  (doppler_max', step) with ``freq_offset_hz`` carrying the difference -- the wipe-off frequency the kernel sees is the same
  integer, so the per-bin result is bit-identical to the unsharded search.
* Tracking units are channels, dealt round-robin.
* Collectives: the sample block is broadcast (or pushed by every rank from its own host copy); results are gathered on rank 0
  as fixed-size records.  The global acquisition winner is picked with the reference's scan order
  (max_to_input_power_statistic, pcps_acquisition.cc:565-596: bins ascending, strict '>'), i.e. the ordering key
  (magnitude descending, Doppler bin ascending, code-phase index ascending), which is what makes sharding invisible.

Works with any initialised torch.distributed backend (nccl on GPUs, gloo in the CPU tests); nothing here touches CUDA.
"""
from dataclasses import dataclass

import numpy as np


def partition(n_units, rank, world):
    """Round-robin ownership: unit u belongs to rank u % world."""
    return list(range(rank, n_units, world))


@dataclass(frozen=True)
class AcqUnit:
    prn_slot: int        # index into the caller's PRN list
    bin_lo: int          # first Doppler bin of the reference grid searched by this unit
    bin_hi: int          # one past the last
    doppler_max: int     # grid parameters to hand to gb200_acq_set_doppler / gb200_acq_cfg for this unit
    doppler_step: int
    freq_offset_hz: int  # added to the caller's own freq_offset_hz (GLONASS FDMA)

    def doppler_hz(self, local_doppler_hz):
        """Map the Doppler the sub-grid search reports back to the reference grid's value."""
        return int(local_doppler_hz) + self.freq_offset_hz


def num_doppler_bins(doppler_max, doppler_step):
    return int(np.ceil(2.0 * doppler_max / doppler_step))  # pcps_acquisition.cc:326


def acquisition_units(n_prn, doppler_max, doppler_step, world):
    """All units of a search, in a fixed global order; unit u is owned by rank u % world."""
    nb = num_doppler_bins(doppler_max, doppler_step)
    splits = 1 if n_prn >= world else min(nb, -(-world // n_prn))
    units = []
    for p in range(n_prn):
        for k in range(splits):
            lo, hi = (nb * k) // splits, (nb * (k + 1)) // splits
            n_sub = hi - lo
            if splits == 1:
                units.append(AcqUnit(p, 0, nb, doppler_max, doppler_step, 0))
                continue
            sub_max = (n_sub * doppler_step) // 2  # ceil(2 sub_max / step) == n_sub
            assert num_doppler_bins(sub_max, doppler_step) == n_sub
            units.append(AcqUnit(p, lo, hi, sub_max, doppler_step, -doppler_max + doppler_step * lo + sub_max))
    return units


ACQ_RECORD = np.dtype([("unit", np.int64), ("prn_slot", np.int64), ("code_phase_index", np.int64), ("doppler_bin", np.int64),
                       ("doppler_hz", np.int64), ("peak_mag", np.float64), ("input_power", np.float64)])


def acq_record(unit_index, unit, code_phase_index, local_bin, local_doppler_hz, peak_mag, input_power):
    r = np.zeros((), ACQ_RECORD)
    r["unit"], r["prn_slot"] = unit_index, unit.prn_slot
    r["code_phase_index"], r["doppler_bin"] = code_phase_index, unit.bin_lo + local_bin
    r["doppler_hz"] = unit.doppler_hz(local_doppler_hz)
    r["peak_mag"], r["input_power"] = peak_mag, input_power
    return r


def merge_acq_winners(records):
    """The winning record of every PRN slot present in `records` (structured array, sorted by prn_slot): one lexsort with the
    reference's ordering key (magnitude desc, Doppler bin asc, code-phase index asc), no per-record Python work."""
    arr = np.asarray(records, dtype=ACQ_RECORD) if not isinstance(records, np.ndarray) else records
    if arr.size == 0:
        return arr
    order = np.lexsort((arr["code_phase_index"], arr["doppler_bin"], -arr["peak_mag"], arr["prn_slot"]))
    srt = arr[order]
    first = np.ones(len(srt), bool)
    first[1:] = srt["prn_slot"][1:] != srt["prn_slot"][:-1]
    return srt[first]


def merge_acq_records(records, n_prn, fft_size, threshold=None):
    """Per PRN: the winning record over its units with the reference's tie-break, and the CFAR statistic
    max / N^4 / input_power in float32 like max_to_input_power_statistic (pcps_acquisition.cc:594)."""
    out = [None] * n_prn
    if len(records) == 0:
        return out
    arr = np.asarray(records, dtype=ACQ_RECORD) if not isinstance(records, np.ndarray) else records
    # one sort with the ordering key (PRN, magnitude desc, Doppler bin asc, code-phase index asc); the first row of every PRN wins
    order = np.lexsort((arr["code_phase_index"], arr["doppler_bin"], -arr["peak_mag"], arr["prn_slot"]))
    srt = arr[order]
    first = np.ones(len(srt), bool)
    first[1:] = srt["prn_slot"][1:] != srt["prn_slot"][:-1]
    nf = np.float32(fft_size) * np.float32(fft_size)
    for best in srt[first]:
        p = int(best["prn_slot"])
        if p < 0 or p >= n_prn:
            continue
        stat = np.float32(best["peak_mag"]) / (nf * nf) / np.float32(best["input_power"]) if best["input_power"] > 0 else np.float32(0)
        d = {k: best[k].item() for k in ACQ_RECORD.names}
        d["test_statistic"] = float(stat)
        if threshold is not None:
            d["decision"] = int(stat > np.float32(threshold))
        out[p] = d
    return out


def gather_records(local_records, dist, dst=0, capacity=None):
    """Gather lists of fixed-size records on `dst` (None elsewhere) with tensor collectives only.  `capacity` = the largest number
    of records any rank can hold (known from the unit list: ceil(n_units / world)); with it the gather is ONE all_gather of
    capacity + 1 records per rank (the first slot carries the count), without it the counts are exchanged first."""
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    arr = np.array(local_records, dtype=ACQ_RECORD) if len(local_records) else np.zeros(0, ACQ_RECORD)
    if capacity is None:
        n = torch.tensor([arr.size], dtype=torch.int64, device=dev)
        counts = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(counts, n)
        capacity = max(int(c.item()) for c in counts)
    assert arr.size <= capacity
    host = np.zeros(capacity + 1, ACQ_RECORD)
    host[0]["unit"] = arr.size
    host[1: 1 + arr.size] = arr
    buf = torch.from_numpy(host.view(np.uint8)).to(dev)
    allbuf = torch.empty(world * buf.numel(), dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(allbuf, buf)
    if rank != dst:
        return None
    full = np.frombuffer(allbuf.cpu().numpy().tobytes(), dtype=ACQ_RECORD).reshape(world, capacity + 1)
    return np.concatenate([full[r, 1: 1 + int(full[r, 0]["unit"])] for r in range(world)])


def gather_channel_outputs(local_channels, local_outputs, n_channels, dist, dst=0):
    """Tracking: rank-local correlator outputs [n_local, epochs, taps] (complex64) -> [n_channels, epochs, taps] on `dst`
    (one all_gather_into_tensor of equal-sized, zero-padded shares)."""
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    local_outputs = np.ascontiguousarray(local_outputs, np.complex64)
    per = -(-n_channels // world)
    shape = (per,) + tuple(local_outputs.shape[1:])
    pad = np.zeros(shape, np.complex64)
    pad[: len(local_channels)] = local_outputs
    t = torch.from_numpy(pad.view(np.float32).reshape(-1)).to(dev, non_blocking=True)
    allt = torch.empty(world * t.numel(), dtype=t.dtype, device=dev)
    dist.all_gather_into_tensor(allt, t)
    if rank != dst:
        return None
    host = allt.cpu().numpy().view(np.complex64).reshape((world,) + shape)
    full = np.zeros((n_channels,) + tuple(local_outputs.shape[1:]), np.complex64)
    for r in range(world):
        chans = partition(n_channels, r, world)
        full[chans] = host[r, : len(chans)]
    return full


def broadcast_samples(tensor, dist, src=0):
    """The conditioned sample block, identical on every rank (ncclBroadcast from the ingest GPU)."""
    dist.broadcast(tensor, src=src)
    return tensor
