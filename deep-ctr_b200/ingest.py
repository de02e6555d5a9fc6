"""Device-side ingestion of the reference's text formats (include/dctr.h dctr_text_*).

The reference parses every sample in Python on the step's critical path (linecache.getline + split, FNN_wnzh.py:224-253;
load_ipinyou_data, ipinyou.py:42-65; get_fxy, baseline.py:190-198).  Here the file's bytes go to the GPU as they are (one
pinned upload), the lines are indexed and parsed there, and the model classes step straight from the resulting device
blocks -- no per-sample host work.  There is no CPU fallback: the host-side loaders of data.py exist for writing fixtures
and for small files, this module is the fast path.
"""
import numpy as np
import torch

from . import _lib as L


def _st(dev):
    return torch.cuda.current_stream(dev).cuda_stream


class DeviceText(object):
    """Raw text resident in HBM + the byte offset of each of its non-empty lines."""

    def __init__(self, data, device="cuda", max_lines=None):
        """data: path, bytes or a uint8 numpy array.  max_lines: capacity of the line index; None = count first (one extra
        indexing pass and a host sync)."""
        L.lib()
        self.dev = torch.device(device)
        if isinstance(data, str):
            data = np.fromfile(data, dtype=np.uint8)
        elif isinstance(data, (bytes, bytearray, memoryview)):
            data = np.frombuffer(data, dtype=np.uint8)
        data = np.ascontiguousarray(data, dtype=np.uint8)
        self.nbytes = int(data.size)
        cap = (self.nbytes + 15) // 16 * 16 + 16
        pinned = torch.zeros(cap, dtype=torch.uint8).pin_memory()
        np.copyto(pinned.numpy()[:self.nbytes], data)         # straight into the pinned block (bytes objects are read-only views)
        self.text = torch.empty(cap, dtype=torch.uint8, device=self.dev)        # torch allocations are >= 256-byte aligned
        self.text.copy_(pinned, non_blocking=True)
        self._pinned = pinned
        self._ws = torch.zeros(max(L.lib().dctr_text_index_workspace(self.nbytes), 8), dtype=torch.uint8, device=self.dev)
        self.n_lines_dev = torch.zeros(1, dtype=torch.int64, device=self.dev)
        if max_lines is None:
            probe = torch.zeros(1, dtype=torch.int64, device=self.dev)
            self._index(probe, 0)
            max_lines = int(self.n_lines_dev.item())
        self.max_lines = int(max_lines)
        self.line_start = torch.zeros(self.max_lines + 1, dtype=torch.int64, device=self.dev)
        self._index(self.line_start, self.max_lines)
        self._n = None

    def _index(self, line_start, max_lines):
        L.call("dctr_text_index_lines", L.ptr(self.text), self.nbytes, L.ptr(line_start), max_lines, L.ptr(self.n_lines_dev),
               L.ptr(self._ws), self._ws.numel(), _st(self.dev))

    @property
    def n_lines(self):
        """number of non-empty lines in the text (host sync on first use); lines past max_lines are not indexed"""
        if self._n is None:
            self._n = int(self.n_lines_dev.item())
        return self._n

    def parse_sparse(self, J, skip=0, pad_id=-1, with_wts=True):
        """`y [skip tokens] id:val ...` -> device tensors (feats int32 [n, J], wts float32 [n, J] or None, y float32 [n],
        status int32 [n]); skip = 0: train.fm.txt (FNN_wnzh.py:248-252), skip = 1: yzx (ipinyou.py:50-59, pad_id = X_dim - 1)."""
        n = self.max_lines
        i32, f32 = dict(dtype=torch.int32, device=self.dev), dict(dtype=torch.float32, device=self.dev)
        feats, y, status = torch.empty(n, J, **i32), torch.empty(n, **f32), torch.empty(n, **i32)
        wts = torch.empty(n, J, **f32) if with_wts else None
        L.call("dctr_text_parse_sparse", L.ptr(self.text), self.nbytes, L.ptr(self.line_start), L.ptr(self.n_lines_dev), n,
               int(skip), int(J), int(pad_id), L.ptr(feats), L.ptr(wts), L.ptr(y), L.ptr(status), _st(self.dev))
        return self._trim(feats, wts, y, status)

    def parse_criteo(self, offsets, max_vals, n_num=13):
        """Criteo indexed TSV -> (ids int32 [n, F], wts float32 [n, F], y float32 [n], status int32 [n]), F = n_num + len(offsets):
        the engines' global-id block (numeric field k = row k weighted by value / max_vals[k]; baseline.py:190-198)."""
        n, n_cat = self.max_lines, len(offsets)
        F = n_num + n_cat
        i32, f32 = dict(dtype=torch.int32, device=self.dev), dict(dtype=torch.float32, device=self.dev)
        ids, wts, y, status = torch.empty(n, F, **i32), torch.empty(n, F, **f32), torch.empty(n, **f32), torch.empty(n, **i32)
        off = torch.as_tensor(np.asarray(offsets, dtype=np.int64)).to(self.dev)
        mv = torch.as_tensor(np.asarray(max_vals, dtype=np.float64)[:max(n_num, 1)]).to(self.dev)
        L.call("dctr_text_parse_criteo", L.ptr(self.text), self.nbytes, L.ptr(self.line_start), L.ptr(self.n_lines_dev), n,
               int(n_num), int(n_cat), L.ptr(off), L.ptr(mv), L.ptr(ids), L.ptr(wts), L.ptr(y), L.ptr(status), _st(self.dev))
        torch.cuda.current_stream(self.dev).synchronize()      # off / mv are temporaries of this call
        return self._trim(ids, wts, y, status)

    def _trim(self, ids, wts, y, status):
        k = min(self.n_lines, self.max_lines)
        return ids[:k], (None if wts is None else wts[:k]), y[:k], status[:k]


def chunk_shuffle_perm(n, size=100000, rng=None):
    """Line order of the reference's `collect` (ipinyou.py:11-20; `size` = 10 000 / 100 000 lines): the file is read in chunks of
    `size` lines and every chunk is shuffled with `np.random.shuffle`.  Returns perm (int64 [n]) with shuffled[k] = lines[perm[k]];
    with the same RandomState it is the reference's order exactly (shuffle draws depend on the chunk length only)."""
    rng = np.random if rng is None else rng
    perm = np.arange(n, dtype=np.int64)
    for s in range(0, n, size):
        rng.shuffle(perm[s:s + size])
    return perm


class LineCursor(object):
    """Which lines the reference's readers hand to each batch, as index arrays into the parsed device blocks.
      wrap=True   get_batch_xy (baseline.py:225-243): read `size` lines; at end of file reopen it and top the batch up from the
                  first line, the next batch continuing from there -- the file is a ring;
      wrap=False  FNN_wnzh.py:292-297 / ipinyou.py:166-172: consecutive full batches, the ragged tail of an epoch is dropped
                  and the next epoch starts at line 0 (next() returns None once at the end of an epoch)."""

    def __init__(self, n_lines, wrap=True):
        self.n, self.wrap, self.pos = int(n_lines), bool(wrap), 0

    def next(self, size):
        if self.n <= 0 or size <= 0:
            return None
        if self.wrap:
            left = self.n - self.pos                       # lines before the end of the file
            take = min(size, left + self.n)                # one reopen per batch: a file shorter than the top-up gives a short batch
            idx = (self.pos + np.arange(take, dtype=np.int64)) % self.n
            self.pos = int((self.pos + take) % self.n)
            return idx
        if self.pos + size > self.n:
            self.pos = 0
            return None
        idx = np.arange(self.pos, self.pos + size, dtype=np.int64)
        self.pos += size
        return idx


def take_lines(perm, *tensors):
    """Apply a line permutation to parsed device blocks (ids, wts, y, ...): a device-side row gather, None passes through."""
    idx = torch.as_tensor(perm, dtype=torch.int64)
    out = []
    for t in tensors:
        out.append(None if t is None else t.index_select(0, idx.to(t.device)))
    return out


def check_status(status, expect=None, max_nnz=None):
    """Raise on lines the parser flagged (the reference's int() / IndexError would have raised on the same lines).
    expect: exact status every line must have (Criteo: 1 + F); max_nnz: sparse lines longer than this were truncated."""
    s = status.cpu().numpy()
    if (s == -2).any():
        raise ValueError("malformed token on line(s) %s" % np.flatnonzero(s == -2)[:5])
    if expect is not None and (s != expect).any():
        raise ValueError("line(s) %s do not have %d fields" % (np.flatnonzero(s != expect)[:5], expect))
    if max_nnz is not None and (s > max_nnz).any():
        raise ValueError("line(s) %s have more than %d features" % (np.flatnonzero(s > max_nnz)[:5], max_nnz))
    return s
