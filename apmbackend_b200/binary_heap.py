"""Min-heap reorder buffer with the reference's method names (binary_heap.js:1-128).
It stays on the host: the heap carries whole TxEntry objects to the db_insert queue and
feeds no statistic.  Order among equal scores is unspecified in both."""
import heapq
import itertools


class BinaryHeap:
    def __init__(self, scoreFunction, content=None):
        self.scoreFunction = scoreFunction
        self._seq = itertools.count()
        self._h = []
        for el in content or []:
            self.push(el)

    @property
    def content(self):
        return [e for _, _, e in self._h]

    def push(self, element):
        heapq.heappush(self._h, (self.scoreFunction(element), next(self._seq), element))

    def pop(self):
        return heapq.heappop(self._h)[2]

    def peek(self):
        return self._h[0][2]

    def size(self):
        return len(self._h)

    def popAllLessOrEqualToScore(self, score):
        out = []
        while self._h and self._h[0][0] <= score:
            out.append(heapq.heappop(self._h)[2])
        return out
