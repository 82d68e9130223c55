"""Host-side synchronisation primitives with the reference's surface
(reference utils.py:4-39): ``TrafficLight.get/switch`` and
``Counter.get/increment/reset``.

As in the reference both ignore their constructor argument: the light starts
False (utils.py:7) and the counter at 0 (utils.py:22).  In the B200 build the
barrier they implement between workers and chief is subsumed by the gradient
exchange inside K3 (one process per GPU); the objects stay so that code
written against the reference keeps working, and ``chief.chief`` still polls
them when it is run as a separate thread."""
import multiprocessing as mp


class TrafficLight:
    """Boolean flip-flop the chief toggles to release the workers."""

    def __init__(self, val=True):
        self._flag = mp.Value("b", False)      # lock comes with the Value

    def get(self):
        with self._flag.get_lock():
            return bool(self._flag.value)

    def switch(self):
        with self._flag.get_lock():
            self._flag.value = not self._flag.value


class Counter:
    """Number of workers that have contributed a gradient since the last chief step."""

    def __init__(self, val=True):
        self._count = mp.Value("i", 0)

    def get(self):
        with self._count.get_lock():
            return int(self._count.value)

    def increment(self):
        with self._count.get_lock():
            self._count.value += 1

    def reset(self):
        with self._count.get_lock():
            self._count.value = 0
