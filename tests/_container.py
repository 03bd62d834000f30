"""Parser of the b200ans blocked container (layout documented in include/b200ans.h)."""
import struct

import numpy as np

MAGIC = 0x4E413242
HEADER = 64


def align(x, a):
    return (x + a - 1) // a * a


def parse(blob):
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    magic, version, method, fidelity, n, block_syms, n_blocks, prelude_bytes, flags, payload_bytes = struct.unpack_from(
        "<IHBBQIIIIQ", blob.tobytes()[:HEADER])
    assert magic == MAGIC and version == 1
    table_at = HEADER + align(prelude_bytes, 4)
    payload_at = align(table_at + 4 * n_blocks, 16)
    sizes = blob[table_at : table_at + 4 * n_blocks].view(np.uint32).copy()
    padded = (sizes.astype(np.uint64) + 15) // 16 * 16
    offsets = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.uint64) if n_blocks else np.zeros(0, np.uint64)
    assert int(padded.sum()) == payload_bytes
    assert payload_at + payload_bytes == blob.size
    blocks = [blob[payload_at + int(o) : payload_at + int(o) + int(s)] for o, s in zip(offsets, sizes)]
    return dict(method=method, fidelity=fidelity, n=n, block_syms=block_syms, n_blocks=n_blocks,
                prelude=blob[HEADER : HEADER + prelude_bytes].copy(), sizes=sizes, offsets=offsets, blocks=blocks,
                payload_at=payload_at, payload_bytes=payload_bytes, pad_ok=_pads_zero(blob, payload_at, offsets, sizes, padded))


def _pads_zero(blob, payload_at, offsets, sizes, padded):
    for o, s, p in zip(offsets, sizes, padded):
        if blob[payload_at + int(o) + int(s) : payload_at + int(o) + int(p)].any():
            return False
    return True


def prelude_equal(ours, theirs):
    """Equal except the reference's uninitialised padding bits in the last 32-bit interp word
    (bits.hpp:148): ours are zero, so `theirs` may only have extra high bits there."""
    ours = np.asarray(ours, dtype=np.uint8)
    theirs = np.asarray(theirs, dtype=np.uint8)
    if ours.size != theirs.size or ours.size < 4:
        return False
    if not np.array_equal(ours[:-4], theirs[:-4]):
        return False
    a = int.from_bytes(ours[-4:].tobytes(), "little")
    b = int.from_bytes(theirs[-4:].tobytes(), "little")
    diff = a ^ b
    if diff == 0:
        return True
    # every differing bit must lie above every bit set in ours (i.e. inside the padding) and be set in theirs
    return (diff & a) == 0 and (diff >> a.bit_length()) << a.bit_length() == diff
