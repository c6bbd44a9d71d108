"""Where the per-fragment ERIs go (reference: ``file_eri_handler.create_dataset(fobj.dname, data=eri)``
at ``src/quemb/molbe/eri_sparse_DF.py:631``, read back by ``get_eri`` ``src/quemb/molbe/helper.py:154-191`` and
``BE.initialize`` ``src/quemb/molbe/mbe.py:1146-1170``).

The reference's container is an HDF5 file (``h5py.File(eri_file, "w")``, ``mbe.py:1226``) with one contiguous float64
dataset ``f{i}`` of shape ``(npair, npair)`` per fragment.  ``open_eri_file`` returns exactly that when h5py is
importable, so the solvers read the file unchanged.  h5py/libhdf5 are absent from the build and GPU boxes of this
repository; there the same interface (``create_dataset`` / ``get`` / ``__getitem__`` / ``close``) is served by a
directory of ``.npy`` files -- same names, same shape, same dtype, memory-mappable -- which is what the tests and the
end-to-end timeline use.  It is a stand-in for the container only; nothing in it is specific to the B200 path.
"""

from __future__ import annotations

import os
import queue
import threading
import time
from concurrent.futures import Future
from pathlib import Path

import numpy as np


class NpyDirFile:
    """``h5py.File``-shaped view of a directory ``<path>.d/`` holding one ``<name>.npy`` per dataset."""

    def __init__(self, path, mode: str = "r"):
        self.path = Path(str(path) + ".d")
        self.mode = mode
        if mode == "w":
            self.path.mkdir(parents=True, exist_ok=True)
            for f in self.path.glob("*.npy"):
                f.unlink()
        elif not self.path.is_dir():
            raise FileNotFoundError(self.path)

    def create_dataset(self, name: str, data: np.ndarray) -> None:
        if self.mode == "r":
            raise OSError("file is open read-only")
        a = np.asarray(data)
        tmp = self.path / f".{name}.tmp.npy"
        with open(tmp, "wb") as fh:  # G is symmetric: C and Fortran order coincide, write without a transposed copy
            np.lib.format.write_array(fh, a if a.flags.c_contiguous else a.T, allow_pickle=False)
        os.replace(tmp, self.path / f"{name}.npy")

    def get(self, name: str):
        f = self.path / f"{name}.npy"
        return np.load(f, mmap_mode="r") if f.exists() else None

    def __getitem__(self, name: str):
        a = self.get(name)
        if a is None:
            raise KeyError(name)
        return a

    def keys(self):
        return sorted(f.stem for f in self.path.glob("*.npy"))

    def close(self) -> None:
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def open_eri_file(path, mode: str = "r"):
    """``h5py.File(path, mode)`` when h5py exists (the reference's container), else :class:`NpyDirFile`."""
    try:
        import h5py  # noqa: PLC0415

        return h5py.File(path, mode)
    except ImportError:
        return NpyDirFile(path, mode)


class AsyncDatasetWriter:
    """Background writer between the pinned output ring of the fragment loop and the container
    (``file_eri_handler.create_dataset(fobj.dname, data=eri)``, ``eri_sparse_DF.py:631``).

    ``submit(name, data)`` returns a ``concurrent.futures.Future`` immediately; one writer thread performs the
    ``create_dataset`` calls in submission order (h5py serialises writers anyway, SURVEY 8b "Threading").  The
    fragment loop (``FragmentERIBuilder.build``) treats a Future returned by its sink as "this pinned buffer is still
    being read" and does not hand the buffer back to the device before it is done, so ``data`` may be a view of the
    pinned ring: no extra host copy of the ``npair^2`` matrix is made, and the disk write of fragment f overlaps the
    kernels and the device->host copy of fragments f+1, f+2."""

    def __init__(self, file_handler, max_pending: int = 4):
        self._fh = file_handler
        self._q: queue.Queue = queue.Queue(maxsize=max_pending)
        self.bytes_written = 0
        self.busy_s = 0.0
        self._thread = threading.Thread(target=self._run, name="qemb-eri-writer", daemon=True)
        self._thread.start()

    def _run(self) -> None:
        while True:
            item = self._q.get()
            if item is None:
                return
            name, data, fut = item
            if not fut.set_running_or_notify_cancel():
                continue
            try:
                t0 = time.perf_counter()
                self._fh.create_dataset(name, data=data)
                self.busy_s += time.perf_counter() - t0
                self.bytes_written += int(getattr(data, "nbytes", 0))
                fut.set_result(name)
            except BaseException as e:  # noqa: BLE001
                fut.set_exception(e)

    def submit(self, name: str, data: np.ndarray) -> Future:
        fut: Future = Future()
        self._q.put((name, data, fut))
        return fut

    def close(self) -> None:
        """Flush: returns when every submitted dataset has been written (errors surface through the futures)."""
        if self._thread.is_alive():
            self._q.put(None)
            self._thread.join()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
