"""No-op timer with the interface `Loss` expects (reference utils/timer.py:19-50 FakeTimer)."""


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
