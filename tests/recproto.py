"""Test-side proto2 wire codec for recode.proto (reference: recode.proto:1-20).

Independent of the product's C++ container code so the two can be checked against each other."""


def _varint(v):
    out = bytearray()
    while v >= 0x80:
        out.append((v & 0x7F) | 0x80)
        v >>= 7
    out.append(v)
    return bytes(out)


def _len_field(num, payload):
    return _varint((num << 3) | 2) + _varint(len(payload)) + payload


def _var_field(num, v):
    return _varint((num << 3) | 0) + _varint(v)


def block_literal(data):
    return {"literal": bytes(data)}


def block_cabac(size, cabac, payload_last_byte=None):
    b = {"size": size, "cabac": bytes(cabac), "length_parity": size & 1}
    if size > 1:
        b["last_byte"] = bytes([payload_last_byte])
    return b


def block_skip(size):
    return {"size": size, "skip_coded": True}


def serialize(blocks, metadata_version=None):
    out = bytearray()
    if metadata_version is not None:
        out += _len_field(1, _len_field(1, metadata_version))
    for b in blocks:
        body = bytearray()
        if "size" in b:
            body += _var_field(1, b["size"])
        if "literal" in b:
            body += _len_field(2, b["literal"])
        if "skip_coded" in b:
            body += _var_field(3, 1 if b["skip_coded"] else 0)
        if "cabac" in b:
            body += _len_field(4, b["cabac"])
        if "length_parity" in b:
            body += _var_field(5, b["length_parity"])
        if "last_byte" in b:
            body += _len_field(6, b["last_byte"])
        out += _len_field(2, bytes(body))
    return bytes(out)


def _read_varint(buf, pos):
    v = 0
    sh = 0
    while True:
        b = buf[pos]
        pos += 1
        v |= (b & 0x7F) << sh
        if not b & 0x80:
            return v, pos
        sh += 7


def _parse_msg(buf):
    pos = 0
    out = []
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        num, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _read_varint(buf, pos)
        elif wt == 2:
            n, pos = _read_varint(buf, pos)
            v = bytes(buf[pos:pos + n])
            pos += n
        else:
            raise ValueError("unexpected wire type %d" % wt)
        out.append((num, v))
    return out


def priors_digest(buf):
    """Metadata field 2 (this library's chunked mode only, not in recode.proto): digest of the prior table, or 0."""
    for num, v in _parse_msg(buf):
        if num == 1:
            for n2, v2 in _parse_msg(v):
                if n2 == 2:
                    return v2
    return 0


def parse(buf):
    """-> (metadata_version or None, [block dict])"""
    version = None
    blocks = []
    for num, v in _parse_msg(buf):
        if num == 1:
            for n2, v2 in _parse_msg(v):
                if n2 == 1:
                    version = v2
        elif num == 2:
            b = {}
            names = {1: "size", 2: "literal", 3: "skip_coded", 4: "cabac", 5: "length_parity", 6: "last_byte",
                     7: "streams", 8: "escaped"}   # 7, 8: this library's chunked mode only, not in recode.proto
            for n2, v2 in _parse_msg(v):
                b[names[n2]] = v2
            blocks.append(b)
    return version, blocks
