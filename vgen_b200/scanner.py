"""ScanConfig / ScanResult, mirroring src/scanner.rs:17-69."""
import collections.abc
import itertools
from dataclasses import dataclass, field
from typing import List, Optional

from .address import AddressFormat, GeneratedAddress


@dataclass
class ScanConfig:
    format: AddressFormat = AddressFormat.P2pkh
    count: int = 1                          # matches wanted; vgen_b200.UNLIMITED (usize::MAX) = no limit, which is what the CLI turns 0 into (src/lib.rs:524)
    threads: Optional[int] = None
    gpu_batch_size: Optional[int] = None
    cpu_batch_size: Optional[int] = None
    start: Optional[int] = None
    end: Optional[int] = None


class MatchList(collections.abc.Sequence):
    """The matches of one scan as a read-only sequence of GeneratedAddress. The library's formatted records
    ("<key hex> <address> <wif>" per line, already copied out of the library) are split into objects on first access, so
    a dense scan with tens of thousands of hits does not pay for objects nobody looks at; len() needs no parsing."""

    def __init__(self, text, count, fmt):
        self._text, self._n, self._fmt, self._items = text, int(count), fmt, None

    def _materialize(self):
        if self._items is None:
            toks = str(self._text, "ascii").split()        # bytes or a memoryview of the caller-owned record buffer
            self._items = list(map(GeneratedAddress, toks[1::3], toks[2::3], toks[0::3], itertools.repeat(self._fmt)))
            self._text = None
        return self._items

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        return self._materialize()[i]

    def __iter__(self):
        return iter(self._materialize())

    def __eq__(self, other):
        return list(self) == list(other)

    def __repr__(self):
        return "MatchList(%r)" % (self._materialize(),)


@dataclass
class ScanResult:
    matches: List[GeneratedAddress] = field(default_factory=list)
    operations: int = 0
    elapsed_secs: float = 0.0

    def rate(self):                         # src/scanner.rs:60-68
        return self.operations / self.elapsed_secs if self.elapsed_secs > 0 else 0.0
