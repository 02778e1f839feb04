"""Wall-clock timer with the reference's interface (cost_model/timer.h:15-130: Start/Stop and
seconds / milli / micro / nano readouts)."""
import time


class Timer:
    def __init__(self):
        self._t0 = self._t1 = None

    def start(self):
        self._t0 = time.perf_counter_ns()
        self._t1 = None
        return self

    def stop(self):
        self._t1 = time.perf_counter_ns()
        return self

    def _ns(self):
        end = self._t1 if self._t1 is not None else time.perf_counter_ns()
        return end - self._t0

    def nanoseconds(self):
        return self._ns()

    def microseconds(self):
        return self._ns() / 1e3

    def milliseconds(self):
        return self._ns() / 1e6

    def seconds(self):
        return self._ns() / 1e9

    __enter__ = start

    def __exit__(self, *exc):
        self.stop()
