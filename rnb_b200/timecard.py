"""Per-request timing records that travel with every request through the pipeline.

Inside an RnB checkout the reference's ``rnb_logging.TimeCard`` / ``TimeCardList`` are used
as-is (the runner creates them; steps only pass them along).  This module provides equivalents
with the same attributes and fork/merge rules (reference ``rnb_logging.py:22-142``) so that the
step classes, the tests and ``bench.py`` also run without the reference tree.
"""
import time
from collections import OrderedDict

try:
    from rnb_logging import TimeCard, TimeCardList  # type: ignore  # noqa: F401
except Exception:  # noqa: BLE001 - not inside an RnB checkout

    class TimeCard:
        """Ordered ``event name -> wall-clock time`` plus the GPUs the request visited."""

        def __init__(self, id):
            self.id = id
            self.sub_id = None
            self.num_parent_timings = None
            self.timings = OrderedDict()
            self.gpus = []

        def record(self, key):
            self.timings[key] = time.time()

        def add_gpu(self, gpu):
            self.gpus.append((gpu,))

        def fork(self, sub_id):
            """Child record for one segment of this request; forks do not nest."""
            if self.sub_id is not None:
                raise Exception('Trying to fork from TimeCard(id %d, sub_id %d).'
                                % (self.id, self.sub_id))
            child = TimeCard(self.id)
            child.sub_id = sub_id
            child.num_parent_timings = len(self.timings)
            child.timings = OrderedDict(self.timings)
            child.gpus = list(self.gpus)
            return child

        @staticmethod
        def merge(time_cards):
            """Join the forks of one request: shared prefix once, later events per sub_id."""
            first = time_cards[0]
            keys = list(first.timings.keys())
            for other in time_cards[1:]:
                if list(other.timings.keys()) != keys:
                    raise Exception('Trying to merge TimeCards with different timing keys. '
                                    '%s != %s' % (keys, list(other.timings.keys())))
                if other.num_parent_timings != first.num_parent_timings:
                    raise Exception('Trying to merge TimeCards that were not forked together. '
                                    '%s != %s' % (first.num_parent_timings,
                                                  other.num_parent_timings))
            ordered = sorted(time_cards, key=lambda tc: tc.sub_id)
            merged = TimeCard(first.id)
            for pos, key in enumerate(keys):
                if pos < first.num_parent_timings:
                    merged.timings[key] = ordered[0].timings[key]
                else:
                    for tc in ordered:
                        merged.timings['%s-%s' % (key, tc.sub_id)] = tc.timings[key]
            for step_gpus in zip(*[tc.gpus for tc in ordered]):
                flat = tuple(g for entry in step_gpus for g in entry)
                merged.gpus.append((flat[0],) if len(set(flat)) == 1 else flat)
            return merged

    class TimeCardList:
        """The records of several requests that were batched into one tensor."""

        def __init__(self, time_cards):
            self.time_cards = time_cards

        def record(self, key):
            for tc in self.time_cards:
                tc.record(key)

        def add_gpu(self, gpu):
            for tc in self.time_cards:
                tc.add_gpu(gpu)

        def fork(self, sub_id):
            raise NotImplementedError('TimeCardLists cannot be forked.')
