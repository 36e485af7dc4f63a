"""Small utilities with the roles of the reference's ``cnn_util.py`` (T/cnn_util.py:26-251):
``roll_numpy_batches``, a reusable ``Barrier`` with abort, and ``ImageProducer`` - a background
thread that keeps a bounded number of input batches staged ahead of the training loop."""
from __future__ import annotations

import threading
from typing import Callable

import numpy as np


def roll_numpy_batches(array, batch_size, shift_ratio):
    """Moves a proportion of batches from start to the end of the array (how workers start at
    different offsets of the same data, T/cnn_util.py:52-67)."""
    num_items = array.shape[0]
    assert num_items % batch_size == 0
    num_batches = num_items // batch_size
    starting_batch = int(num_batches * shift_ratio)
    starting_item = starting_batch * batch_size
    return np.roll(array, -starting_item, axis=0)


class Barrier:
    """Reusable barrier for `parties` threads; `abort` releases everybody with an error."""

    def __init__(self, parties: int):
        self.parties = parties
        self._cond = threading.Condition()
        self._count = 0
        self._generation = 0
        self._aborted = False

    def wait(self) -> None:
        with self._cond:
            if self._aborted:
                raise RuntimeError("barrier aborted")
            gen = self._generation
            self._count += 1
            if self._count == self.parties:
                self._generation += 1
                self._count = 0
                self._cond.notify_all()
                return
            while gen == self._generation and not self._aborted:  # spurious wake-ups
                self._cond.wait()
            if self._aborted:
                raise RuntimeError("barrier aborted")

    def abort(self) -> None:
        with self._cond:
            self._aborted = True
            self._cond.notify_all()


class ImageProducer:
    """Runs `produce()` on a background thread so that between 0 and 2*batch_group_size batches
    are staged ahead of the consumer (T/cnn_util.py:118-198)."""

    def __init__(self, produce: Callable[[], object], batch_group_size: int = 1):
        self.produce = produce
        self.batch_group_size = max(int(batch_group_size), 1)
        self.capacity = 2 * self.batch_group_size
        self._items = []
        self._cond = threading.Condition()
        self._done = False
        self._error = None
        self._thread = None

    def start(self) -> None:
        self._thread = threading.Thread(target=self._loop, daemon=True)
        self._thread.start()

    def _loop(self) -> None:
        try:
            while True:
                with self._cond:
                    while len(self._items) + self.batch_group_size > self.capacity and \
                            not self._done:
                        self._cond.wait()
                    if self._done:
                        return
                group = [self.produce() for _ in range(self.batch_group_size)]
                with self._cond:
                    self._items.extend(group)
                    self._cond.notify_all()
        except BaseException as e:  # a failing producer must not leave the consumer waiting
            with self._cond:
                self._error = e
                self._done = True
                self._cond.notify_all()

    def staged(self) -> int:
        with self._cond:
            return len(self._items)

    def get(self):
        """Next staged batch (notify_image_consumption of the reference is implicit)."""
        with self._cond:
            while not self._items and not self._done:
                self._cond.wait()
            if not self._items:
                if self._error is not None:
                    raise RuntimeError("input pipeline failed: %r" % (self._error,)) \
                        from self._error
                raise RuntimeError("producer stopped")
            item = self._items.pop(0)
            self._cond.notify_all()
            return item

    def done(self) -> None:
        with self._cond:
            self._done = True
            self._cond.notify_all()
        if self._thread is not None:
            self._thread.join(timeout=5)


class BaseClusterManager:
    """Cluster description derived from ``--worker_hosts`` / ``--ps_hosts`` / ``--job_name`` /
    ``--task_index`` (T/cnn_util.py:201-251). On this runtime a worker is a rank of the
    process group and parameter servers have no process of their own, so the manager is a
    plain description: who the workers are, where the rendezvous is, what this process is."""

    def __init__(self, params):
        self.worker_hosts = [h for h in (params.worker_hosts or "").split(",") if h]
        self.ps_hosts = [h for h in (params.ps_hosts or "").split(",") if h]
        self.job_name = params.job_name or ""
        self.task_index = int(params.task_index or 0)

    def get_target(self):
        """Rendezvous address (first worker), the counterpart of the TF session target."""
        return self.worker_hosts[0] if self.worker_hosts else ""

    def join_server(self):
        """A TF parameter server blocks here forever; there is nothing to serve."""
        return None

    def get_cluster_spec(self):
        return {"worker": list(self.worker_hosts), "ps": list(self.ps_hosts)}

    def num_workers(self):
        return len(self.worker_hosts)

    def num_ps(self):
        return len(self.ps_hosts)


class GrpcClusterManager(BaseClusterManager):
    """Name kept for code written against the reference; no gRPC server is started."""
