"""provider::resolve / build_pattern / build_exact_pattern (src/provider.rs:12-63) over the library's embedded puzzle table."""
import ctypes as C
from dataclasses import dataclass
from typing import Optional, Tuple

from . import _cabi
from .address import AddressFormat


@dataclass
class ProviderResult:               # src/provider.rs:6-10
    address: str
    format: AddressFormat
    key_range: Optional[Tuple[int, int]]


def resolve(pattern):
    """None when `pattern` is not a provider reference; raises ValueError when the provider cannot resolve it."""
    addr = C.create_string_buffer(128)
    err = C.create_string_buffer(256)
    fmt, has = C.c_int(0), C.c_int(0)
    lo, hi = C.create_string_buffer(32), C.create_string_buffer(32)
    rc = _cabi.lib().vgen_provider_resolve(pattern.encode(), addr, 128, C.byref(fmt), lo, hi, C.byref(has), err, 256)
    if rc == 0:
        return None
    if rc < 0:
        raise ValueError(err.value.decode())
    rng = (int.from_bytes(lo.raw, "big"), int.from_bytes(hi.raw, "big")) if has.value else None
    return ProviderResult(addr.value.decode(), AddressFormat(fmt.value), rng)


def _pattern(result, n):
    buf = C.create_string_buffer(256)
    _cabi.lib().vgen_provider_pattern(result.address.encode(), n, buf, 256)
    return buf.value.decode()


def build_pattern(result, prefix_length):
    return _pattern(result, max(1, int(prefix_length)))


def build_exact_pattern(result):
    return _pattern(result, 0)
