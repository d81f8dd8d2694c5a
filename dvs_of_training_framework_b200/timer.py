"""Timer groups with the interface `Loss` and `train` expect (reference utils/timer.py):
`FakeTimer` (:19-50, the no-op default of `Loss`) and `SynchronizedWallClockTimer` (:52-132, what
`--timers` selects): named wall-clock phases, the device drained at every start/stop, one line per
training iteration printed by rank 0."""
import time

import torch


def get_rank():
    """utils/timer.py:8-11."""
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        return torch.distributed.get_rank()
    return 0


class FakeTimer:
    class Timer:
        def __init__(self, name):
            self.name = name

        def start(self):
            pass

        def stop(self):
            pass

        def reset(self):
            pass

        def elapsed(self, reset=True):
            return 0

    def __init__(self):
        self.timers = {}

    def __call__(self, name):
        return self.timers.setdefault(name, self.Timer(name))

    @staticmethod
    def memory_usage():
        return ''

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False):
        pass


def _drain():
    if torch.cuda.is_available():
        torch.cuda.synchronize()


class SynchronizedWallClockTimer(FakeTimer):
    """utils/timer.py:52-132.  Same phase names, same `rank=R time (ms) | name: x.xx | ...` line."""

    class Timer:
        def __init__(self, name):
            self.name_, self.elapsed_, self.started_ = name, 0.0, False
            self.start_time = time.time()

        def start(self):
            assert not self.started_, 'timer has already been started'
            _drain()
            self.start_time, self.started_ = time.time(), True

        def stop(self):
            assert self.started_, 'timer is not started'
            _drain()
            self.elapsed_ += time.time() - self.start_time
            self.started_ = False

        def reset(self):
            self.elapsed_, self.started_ = 0.0, False

        def elapsed(self, reset=True):
            running = self.started_
            if running:
                self.stop()
            total = self.elapsed_
            if reset:
                self.reset()
            if running:
                self.start()
            return total

    @staticmethod
    def memory_usage():
        gb = 1024 ** 3
        return (f' | mem_allocated: {torch.cuda.memory_allocated() / gb:.4f} GB'
                f' | max_mem_allocated: {torch.cuda.max_memory_allocated() / gb:.4f} GB'
                f' | cache_allocated: {torch.cuda.memory_reserved() / gb:.4f} GB'
                f' | max_cache_allocated: {torch.cuda.max_memory_reserved() / gb:.4f} GB')

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False):
        assert normalizer > 0.0
        parts = [f'rank={get_rank()} time (ms)']
        parts += [f'{n}: {self.timers[n].elapsed(reset=reset) * 1000.0 / normalizer:.2f}'
                  for n in names if n in self.timers]
        if get_rank() == 0:
            print(' | '.join(parts))
