"""Wav emission off the synthesis thread: a background writer fed with int16 sample buffers.

Stands behind the reference's synchronous per-utterance ``scipy.io.wavfile.write(path, sampling_rate, wav)`` loop
(utils/tools.py:272-274, examples_gen.py:108-109).  The files are byte-identical to what scipy writes for a 1-D int16 array
(44-byte RIFF/WAVE header, PCM format 1, mono, little endian), so downstream tools cannot tell the difference
(tests/test_host_logic.py compares the bytes).  File writes release the GIL, so a small pool of Python threads overlaps them
with the next batch's synthesis; ``submit`` copies nothing when the caller hands over ownership of the buffer.
"""
import os
import queue
import struct
import threading

import numpy as np


def wav_bytes_header(n_samples, rate, channels=1, bits=16):
    """The 44-byte header scipy.io.wavfile.write emits for integer PCM data."""
    block = channels * bits // 8
    data = n_samples * block
    return (b"RIFF" + struct.pack("<I", 36 + data) + b"WAVE" + b"fmt " + struct.pack("<IHHIIHH", 16, 1, channels, rate, rate * block, block, bits)
            + b"data" + struct.pack("<I", data))


def write_wav(path, rate, samples):
    """Synchronous single-file write: ``scipy.io.wavfile.write`` for a 1-D int16 array."""
    a = np.ascontiguousarray(samples)
    if a.dtype != np.int16 or a.ndim != 1:
        raise ValueError("write_wav expects a 1-D int16 array, got %s %s" % (a.dtype, a.shape))
    with open(path, "wb") as f:
        f.write(wav_bytes_header(a.shape[0], int(rate)))
        f.write(a.data)


class WavWriter:
    """Background writer: ``submit(path, rate, samples)`` returns at once; ``close()`` (or leaving the ``with`` block) waits for
    the queue to drain and re-raises the first I/O error.  ``samples`` may be a view into a pinned buffer that the caller will
    overwrite — pass ``copy=True`` (default) then; ``copy=False`` hands the array over."""

    def __init__(self, n_threads=4, max_pending=4096):
        self._q = queue.Queue(maxsize=max_pending)
        self._err = None
        self.files = 0
        self.bytes = 0
        self._lock = threading.Lock()
        self._threads = [threading.Thread(target=self._run, daemon=True) for _ in range(max(1, n_threads))]
        for t in self._threads:
            t.start()

    def _run(self):
        while True:
            item = self._q.get()
            if item is None:
                self._q.task_done()
                return
            path, rate, a = item
            try:
                write_wav(path, rate, a)
                with self._lock:
                    self.files += 1
                    self.bytes += 44 + a.nbytes
            except Exception as e:                                  # noqa: BLE001 - reported by close()
                with self._lock:
                    if self._err is None:
                        self._err = e
            finally:
                self._q.task_done()

    def submit(self, path, rate, samples, copy=True):
        if self._err is not None:
            raise self._err
        a = np.asarray(samples)
        if a.dtype != np.int16 or a.ndim != 1:
            raise ValueError("WavWriter expects 1-D int16 arrays, got %s %s" % (a.dtype, a.shape))
        d = os.path.dirname(str(path))
        if d and not os.path.isdir(d):
            os.makedirs(d, exist_ok=True)
        self._q.put((str(path), int(rate), a.copy() if copy else a))

    def flush(self):
        self._q.join()
        if self._err is not None:
            raise self._err

    def close(self):
        self._q.join()
        for _ in self._threads:
            self._q.put(None)
        for t in self._threads:
            t.join()
        self._threads = []
        if self._err is not None:
            raise self._err

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False
