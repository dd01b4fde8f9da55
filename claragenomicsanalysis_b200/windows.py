"""Synthetic ONT-like POA windows and the flat window-set container every entry point takes.

The reference benchmarks read cudapoa/data/sample-windows.txt, which is not part of the snapshot; these
deterministic windows replace it.  Window i is a pure function of (seed0 + i) and the shape arguments.
"""
import numpy as np

SEED0 = 20261104
_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


class WindowSet:
    """Windows as flat buffers: ``bases`` (uint8, reads of all windows back to back), ``read_lengths`` (int32, one
    per read), ``window_n_reads`` (int32, one per window) and optional ``weights`` (int8, parallel to bases)."""

    def __init__(self, bases, read_lengths, window_n_reads, weights=None):
        self.bases = np.ascontiguousarray(bases, dtype=np.uint8)
        self.read_lengths = np.ascontiguousarray(read_lengths, dtype=np.int32)
        self.window_n_reads = np.ascontiguousarray(window_n_reads, dtype=np.int32)
        self.weights = None if weights is None else np.ascontiguousarray(weights, dtype=np.int8)
        self._read_off = np.concatenate([[0], np.cumsum(self.window_n_reads, dtype=np.int64)])
        self._base_off_read = np.concatenate([[0], np.cumsum(self.read_lengths, dtype=np.int64)])

    def __len__(self):
        return int(self.window_n_reads.shape[0])

    @property
    def n_reads(self):
        return int(self.read_lengths.shape[0])

    @property
    def n_bases(self):
        return int(self.bases.shape[0])

    @property
    def max_read_length(self):
        return int(self.read_lengths.max()) if self.n_reads else 0

    @property
    def max_reads_per_window(self):
        return int(self.window_n_reads.max()) if len(self) else 0

    def window(self, i):
        """Reads of window i as a list of str."""
        out = []
        for r in range(int(self._read_off[i]), int(self._read_off[i + 1])):
            a, b = int(self._base_off_read[r]), int(self._base_off_read[r + 1])
            out.append(self.bases[a:b].tobytes().decode("ascii"))
        return out

    def window_weights(self, i):
        if self.weights is None:
            return None
        out = []
        for r in range(int(self._read_off[i]), int(self._read_off[i + 1])):
            a, b = int(self._base_off_read[r]), int(self._base_off_read[r + 1])
            out.append(self.weights[a:b].copy())
        return out

    def slice(self, start, stop):
        """Windows [start, stop) as a new WindowSet (views where possible)."""
        start = max(0, min(start, len(self)))
        stop = max(start, min(stop, len(self)))
        r0, r1 = int(self._read_off[start]), int(self._read_off[stop])
        b0, b1 = int(self._base_off_read[r0]), int(self._base_off_read[r1])
        w = None if self.weights is None else self.weights[b0:b1]
        return WindowSet(self.bases[b0:b1], self.read_lengths[r0:r1], self.window_n_reads[start:stop], w)

    def tile(self, n_windows):
        """Repeat the set cyclically up to n_windows windows, the way the reference tiles its input file
        (cudapoa/include/.../utils.hpp:67-86)."""
        if n_windows <= len(self):
            return self.slice(0, n_windows)
        reps, rem = divmod(n_windows, len(self))
        parts = [self] * reps + ([self.slice(0, rem)] if rem else [])
        return WindowSet.concat(parts)

    @staticmethod
    def concat(parts):
        w = None
        if all(p.weights is not None for p in parts):
            w = np.concatenate([p.weights for p in parts])
        return WindowSet(np.concatenate([p.bases for p in parts]),
                         np.concatenate([p.read_lengths for p in parts]),
                         np.concatenate([p.window_n_reads for p in parts]), w)

    @staticmethod
    def from_lists(windows, weights=None):
        """Build from a list of windows, each a list of str/bytes."""
        bases, lens, counts, wts = [], [], [], []
        for wi, win in enumerate(windows):
            counts.append(len(win))
            for ri, seq in enumerate(win):
                b = seq.encode("ascii") if isinstance(seq, str) else bytes(seq)
                bases.append(np.frombuffer(b, dtype=np.uint8))
                lens.append(len(b))
                if weights is not None:
                    wts.append(np.asarray(weights[wi][ri], dtype=np.int8))
        cat = np.concatenate(bases) if bases else np.zeros(0, dtype=np.uint8)
        w = (np.concatenate(wts) if wts else np.zeros(0, dtype=np.int8)) if weights is not None else None
        return WindowSet(cat, np.asarray(lens, dtype=np.int32), np.asarray(counts, dtype=np.int32), w)


def _mutate(rng, template, sub, ins, dele):
    """One noisy copy of `template` (values 0..3): i.i.d. per-position substitution / insertion / deletion."""
    n = template.shape[0]
    u = rng.random(n)
    shift = rng.integers(1, 4, n, dtype=np.int64)
    v = rng.random(n)
    ins_base = rng.integers(0, 4, n, dtype=np.int64)
    keep = u >= dele
    is_sub = keep & (u < dele + sub)
    b = template.astype(np.int64).copy()
    b[is_sub] = (b[is_sub] + shift[is_sub]) % 4
    insm = v < ins
    counts = keep.astype(np.int64) + insm.astype(np.int64)
    start = np.cumsum(counts) - counts
    out = np.empty(int(counts.sum()), dtype=np.int64)
    out[start[keep]] = b[keep]
    out[(start + keep)[insm]] = ins_base[insm]
    if out.shape[0] == 0:  # never hand the engine an empty read
        out = b[:1]
    return out


def generate_window(index, template_length=500, n_reads=30, seed0=SEED0, backbone_error=0.05, read_error=0.10,
                    with_weights=False, length_choices=None, length_probs=None, n_reads_range=None):
    """Window `index`: template uniform over ACGT; read 0 = template with `backbone_error` errors; reads 1.. =
    template with `read_error` errors, split 30 % substitutions / 30 % insertions / 40 % deletions."""
    rng = np.random.default_rng(seed0 + int(index))
    lt = template_length
    if length_choices is not None:
        lt = int(rng.choice(np.asarray(length_choices), p=length_probs))
    elif isinstance(template_length, (tuple, list)):
        lt = int(rng.integers(template_length[0], template_length[1] + 1))
    s = n_reads
    if n_reads_range is not None:
        s = int(rng.integers(n_reads_range[0], n_reads_range[1] + 1))
    template = rng.integers(0, 4, lt, dtype=np.int64)
    reads = []
    for r in range(s):
        e = backbone_error if r == 0 else read_error
        reads.append(_ACGT[_mutate(rng, template, 0.3 * e, 0.3 * e, 0.4 * e)])
    weights = None
    if with_weights:
        weights = [rng.integers(0, 41, rd.shape[0], dtype=np.int64).astype(np.int8) for rd in reads]
    return reads, weights


def generate_windows(n_windows, start_index=0, **kwargs):
    """`n_windows` consecutive windows starting at `start_index` as a WindowSet."""
    bases, lens, counts, wts = [], [], [], []
    with_weights = kwargs.get("with_weights", False)
    for i in range(start_index, start_index + n_windows):
        reads, weights = generate_window(i, **kwargs)
        counts.append(len(reads))
        for rd in reads:
            bases.append(rd)
            lens.append(rd.shape[0])
        if with_weights:
            wts.extend(weights)
    return WindowSet(np.concatenate(bases) if bases else np.zeros(0, np.uint8), np.asarray(lens, np.int32),
                     np.asarray(counts, np.int32), np.concatenate(wts) if with_weights else None)


# Named shapes of BASELINE.json's configs.
def config_windows(name, n_windows, start_index=0, with_weights=False):
    if name in ("cfg2", "cfg3", "cfg5a"):
        return generate_windows(n_windows, start_index, template_length=500, n_reads=30, with_weights=with_weights)
    if name == "cfg4":
        return generate_windows(n_windows, start_index, template_length=(2000, 5000), n_reads=100,
                                with_weights=with_weights)
    if name == "cfg5b":
        return generate_windows(n_windows, start_index, length_choices=[250, 500, 1000, 2000],
                                length_probs=[0.2, 0.5, 0.2, 0.1], n_reads_range=(10, 60),
                                with_weights=with_weights)
    raise ValueError("unknown config " + name)


def _config_chunk(args):
    name, n, start, with_weights = args
    ws = config_windows(name, n, start_index=start, with_weights=with_weights)
    return ws.bases, ws.read_lengths, ws.window_n_reads, ws.weights


def config_windows_parallel(name, n_windows, start_index=0, with_weights=False, processes=None):
    """config_windows over a process pool (window i depends on seed0 + i only, so the result is identical).  Call it
    before CUDA is initialised in the process: the pool forks."""
    import multiprocessing as mp
    import os

    if processes is None:
        try:
            processes = len(os.sched_getaffinity(0))
        except AttributeError:
            processes = os.cpu_count() or 1
    processes = max(1, min(processes, 32, n_windows // 64))
    if processes <= 1:
        return config_windows(name, n_windows, start_index=start_index, with_weights=with_weights)
    step = (n_windows + 4 * processes - 1) // (4 * processes)
    jobs = [(name, min(step, n_windows - at), start_index + at, with_weights) for at in range(0, n_windows, step)]
    with mp.get_context("fork").Pool(processes) as pool:
        parts = pool.map(_config_chunk, jobs)
    return WindowSet.concat([WindowSet(*p) for p in parts])
