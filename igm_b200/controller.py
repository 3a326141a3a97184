"""A parallel controller that turns IGM's one-structure-per-task M-step into ONE GPU launch.

``ModelingStep`` hands ``controller.map(task, range(population_size))`` one task per structure
(igm/core/step.py:259-274, igm/steps/ModelingStep.py:97); every task builds its Model and calls
``model.optimize(cfg)``.  Registered under a name in ``igm.parallel.controller_class``
(igm/parallel/__init__.py:9-20), ``B200Controller`` runs the unchanged tasks on host threads and
parks their ``optimize`` calls in a batcher; once every running task is parked (or ``batch_size``
are), one of them launches ``optimize_population`` for all and everybody returns with its own
result.  Per-structure ``.hms`` / ``.ready`` files and the restart logic are untouched, since the
tasks themselves do not change.  With any other controller ``igm_b200.kernel.optimize`` simply runs
a batch of one.
"""
import threading

from . import kernel

_tls = threading.local()


class PopulationBatcher(object):
    """Rendezvous of concurrent ``optimize(model, cfg)`` calls."""

    def __init__(self, batch_size=None, run=None):
        self._cv = threading.Condition()
        self._live = 0                 # workers that still have tasks (parked or not)
        self._parked = []              # [model, cfg, slot] waiting for a launch
        self._batch_size = batch_size
        self._run = run or kernel.optimize_population
        self.batches = []              # sizes of the launches made (diagnostics / tests)

    # -- worker bookkeeping: a worker counts as live from the start of map() until it runs out of tasks --
    def set_workers(self, n):
        with self._cv:
            self._live = n

    def worker_done(self):
        with self._cv:
            self._live -= 1
            self._cv.notify_all()      # the others may now be "everybody"

    def _ready(self):
        n = len(self._parked)
        return n > 0 and (n >= self._live or (self._batch_size is not None and n >= self._batch_size))

    def optimize(self, model, cfg):
        slot = {"done": False, "info": None, "error": None}
        with self._cv:
            self._parked.append((model, cfg, slot))
            while not slot["done"]:
                if self._ready():
                    batch, self._parked = self._parked, []
                    self._launch(batch)        # releases the lock while the GPU works
                    continue
                self._cv.wait()
        if slot["error"] is not None:
            raise slot["error"]
        return slot["info"]

    def _launch(self, batch):
        # called with the lock held; the batch has been taken out of the parking list
        self._live -= len(batch)               # parked callers do not count while their launch runs
        self._cv.release()
        try:
            try:
                infos = self._run([m for m, _, _ in batch], batch[0][1])
                err = None
            except Exception as e:             # noqa: BLE001 -- handed to every caller of the batch
                infos, err = [None] * len(batch), e
        finally:
            self._cv.acquire()
        self._live += len(batch)
        self.batches.append(len(batch))
        for (_, _, slot), info in zip(batch, infos):
            slot["info"], slot["error"], slot["done"] = info, err, True
        self._cv.notify_all()


def current_batcher():
    return getattr(_tls, "batcher", None)


class B200Controller(object):
    """Drop-in for the classes in igm/parallel (same ``setup/map/reduce/map_reduce/teardown``)."""

    def __init__(self, max_threads=256, batch_size=None):
        self.max_threads = int(max_threads)
        self.batch_size = batch_size

    def setup(self):
        pass

    def teardown(self):
        pass

    def map(self, parallel_task, args):
        args = list(args)
        if not args:
            return []
        batcher = PopulationBatcher(self.batch_size)
        self.last_batcher = batcher
        n_workers = min(len(args), self.max_threads)
        batcher.set_workers(n_workers)
        results, errors = [None] * len(args), []
        lock = threading.Lock()
        cursor = [0]

        def worker():
            _tls.batcher = batcher
            try:
                while True:
                    with lock:
                        k = cursor[0]
                        cursor[0] += 1
                    if k >= len(args) or errors:
                        return
                    try:
                        results[k] = parallel_task(args[k])
                    except BaseException as e:      # noqa: BLE001 -- re-raised in the caller's thread below
                        errors.append(e)
                        return
            finally:
                _tls.batcher = None
                batcher.worker_done()

        threads = [threading.Thread(target=worker, daemon=True) for _ in range(n_workers)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return results

    def reduce(self, reduce_task, outs):
        return reduce_task(outs)

    def map_reduce(self, parallel_task, reduce_task, args):
        return self.reduce(reduce_task, self.map(parallel_task, args))
