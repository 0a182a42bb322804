"""ScanConfig / ScanResult, mirroring src/scanner.rs:17-69."""
from dataclasses import dataclass, field
from typing import List, Optional

from .address import AddressFormat, GeneratedAddress


@dataclass
class ScanConfig:
    format: AddressFormat = AddressFormat.P2pkh
    count: int = 1                          # 0 = unlimited (src/lib.rs:524 maps 0 -> usize::MAX)
    threads: Optional[int] = None
    gpu_batch_size: Optional[int] = None
    cpu_batch_size: Optional[int] = None
    start: Optional[int] = None
    end: Optional[int] = None


@dataclass
class ScanResult:
    matches: List[GeneratedAddress] = field(default_factory=list)
    operations: int = 0
    elapsed_secs: float = 0.0

    def rate(self):                         # src/scanner.rs:60-68
        return self.operations / self.elapsed_secs if self.elapsed_secs > 0 else 0.0
