"""Processor plumbing: the same calling conventions as madmom/processors.py so that chains written for the
reference run unchanged on the GPU classes.

  Processor            process(data, **kwargs); __call__ -> process; dump/load = pickle
                       (madmom/processors.py:27-114)
  _process             Processor instances receive **kwargs, bare callables (np.hstack, superflux, ...) receive
                       only the data, None passes the data through (madmom/processors.py:247-285)
  SequentialProcessor  lists / tuples nest automatically (madmom/processors.py:288-418)
  ParallelProcessor    same data to every processor, returns a list (madmom/processors.py:422-479).  On the
                       GPU every branch is an asynchronous kernel launch on the same stream, so `num_threads`
                       is accepted and ignored.
"""
import pickle
from collections.abc import MutableSequence


class Processor(object):
    """Abstract base class of everything that processes data."""

    @classmethod
    def load(cls, infile):
        """Un-pickle a Processor from a file name or file handle (madmom/processors.py:36-62)."""
        if hasattr(infile, "read"):
            return pickle.load(infile)
        with open(infile, "rb") as f:
            return pickle.load(f)

    def dump(self, outfile):
        """Pickle the Processor to a file name or file handle (madmom/processors.py:64-88)."""
        if hasattr(outfile, "write"):
            pickle.dump(self, outfile, protocol=2)
        else:
            with open(outfile, "wb") as f:
                pickle.dump(self, f, protocol=2)

    def process(self, data, **kwargs):
        raise NotImplementedError("Must be implemented by subclass.")

    def __call__(self, *args, **kwargs):
        return self.process(*args, **kwargs)


def _process(process_tuple):
    """(processor-or-callable, data[, output], kwargs) -> processed data."""
    proc = process_tuple[0]
    if proc is None:
        return process_tuple[1]
    if isinstance(proc, Processor):
        return proc(*process_tuple[1:-1], **process_tuple[-1])
    return proc(*process_tuple[1:-1])


class SequentialProcessor(MutableSequence, Processor):
    """Feed the output of each processor into the next one."""

    def __init__(self, processors):
        self.processors = []
        for p in processors:
            self.processors.append(SequentialProcessor(p) if isinstance(p, (list, tuple)) else p)

    def __getitem__(self, index):
        return self.processors[index]

    def __setitem__(self, index, processor):
        self.processors[index] = processor

    def __delitem__(self, index):
        del self.processors[index]

    def __len__(self):
        return len(self.processors)

    def insert(self, index, processor):
        self.processors.insert(index, processor)

    def append(self, other):
        self.processors.append(other)

    def extend(self, other):
        self.processors.extend(other)

    def process(self, data, **kwargs):
        for p in self.processors:
            data = _process((p, data, kwargs))
        return data


class ParallelProcessor(SequentialProcessor):
    """Feed the same data to every processor; returns the list of results."""

    def __init__(self, processors, num_threads=None):
        super(ParallelProcessor, self).__init__(processors)
        self.num_threads = 1 if num_threads is None else num_threads

    def process(self, data, **kwargs):
        return [_process((p, data, kwargs)) for p in self.processors]
