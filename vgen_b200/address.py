"""AddressFormat / GeneratedAddress, mirroring src/address.rs:11-72 of the reference. String encoding of a hit
(address from digest, WIF from key) is done by the library's host layer (vgen_format_address / vgen_format_wif)."""
import ctypes as C
import typing
import enum
from dataclasses import dataclass

from . import _cabi


class AddressFormat(enum.IntEnum):
    # values = enum vgen_format (include/vgen_cuda.h)
    P2pkh = 0
    P2pkhUncompressed = 1
    P2wpkh = 2
    P2shP2wpkh = 3
    P2tr = 4
    Ethereum = 5

    @staticmethod
    def all():                      # src/address.rs:27-36
        return [AddressFormat.P2pkh, AddressFormat.P2pkhUncompressed, AddressFormat.P2wpkh,
                AddressFormat.P2shP2wpkh, AddressFormat.P2tr, AddressFormat.Ethereum]

    def charset_name(self):         # src/address.rs:39-45
        if self in (AddressFormat.P2pkh, AddressFormat.P2pkhUncompressed, AddressFormat.P2shP2wpkh):
            return "Base58"
        return "Bech32" if self in (AddressFormat.P2wpkh, AddressFormat.P2tr) else "Hex"

    def __str__(self):              # Display, src/address.rs:48-59
        return {0: "P2PKH", 1: "P2PKH (Uncompressed)", 2: "P2WPKH", 3: "P2SH-P2WPKH", 4: "P2TR", 5: "Ethereum"}[int(self)]

    @staticmethod
    def parse(s):                   # CLI value names (src/lib.rs:213-242)
        if isinstance(s, AddressFormat):
            return s
        if isinstance(s, int):
            return AddressFormat(s)
        table = {"p2pkh": 0, "p2pkh-uncompressed": 1, "p2wpkh": 2, "p2sh-p2wpkh": 3, "p2tr": 4, "ethereum": 5}
        return AddressFormat(table[s.lower()])


class GeneratedAddress(typing.NamedTuple):      # src/address.rs:62-72 (a tuple type: dense scans build tens of thousands per call)
    address: str
    wif: str
    hex: str
    format: AddressFormat


def address_from_digest(fmt, digest):
    buf = C.create_string_buffer(128)
    n = _cabi.lib().vgen_format_address(int(fmt), bytes(digest).ljust(32, b"\0"), buf, 128)
    if n < 0:
        raise _cabi.VgenError(n, "vgen_format_address")
    return buf.value.decode()


def wif_from_key(fmt, key_be):
    buf = C.create_string_buffer(128)
    n = _cabi.lib().vgen_format_wif(int(fmt), bytes(key_be), buf, 128)
    if n < 0:
        raise _cabi.VgenError(n, "vgen_format_wif")
    return buf.value.decode()
