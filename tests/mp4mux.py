"""Minimal ISO base media (mp4) muxer for tests: wraps an Annex-B H.264 elementary stream into
ftyp + moov + mdat with an avcC box and length-prefixed NAL units, the way cameras write it (the
reference's documented input is a GoPro .MP4, README.md:47-49, demuxed by libavformat).  One sample
per access unit; SPS / PPS go into avcC and are dropped from the samples.  Test infrastructure only."""
import struct


def split_annexb(data):
    """-> list of NAL unit byte strings (header byte first, emulation-prevention bytes kept)."""
    nals = []
    i = data.find(b"\x00\x00\x01")
    while i >= 0:
        start = i + 3
        j = data.find(b"\x00\x00\x01", start)
        end = len(data) if j < 0 else j
        nal = data[start:end]
        if j >= 0:
            nal = nal.rstrip(b"\x00") if nal[-1:] == b"\x00" else nal   # the zero of a 4-byte start code
        nals.append(nal)
        i = j
    return nals


def box(kind, *payload):
    body = b"".join(payload)
    return struct.pack(">I4s", 8 + len(body), kind) + body


def full(kind, version, flags, *payload):
    return box(kind, struct.pack(">I", version << 24 | flags), *payload)


def mux(data, width, height, nal_len=4, timescale=30, chunk=3, co64=False):
    nals = split_annexb(data)
    sps = [n for n in nals if n[0] & 31 == 7]
    pps = [n for n in nals if n[0] & 31 == 8]
    assert sps and pps
    # access units: a new one starts at every slice with first_mb_in_slice == 0 (ue(0) = first bit set)
    samples, cur = [], b""
    for n in nals:
        t = n[0] & 31
        if t in (7, 8):
            continue
        if t in (1, 5) and n[1] & 0x80 and cur:
            samples.append(cur)
            cur = b""
        cur += len(n).to_bytes(nal_len, "big") + n
    if cur:
        samples.append(cur)
    avcc = box(b"avcC", bytes([1, sps[0][1], sps[0][2], sps[0][3], 0xFC | (nal_len - 1), 0xE0 | len(sps)]),
               b"".join(struct.pack(">H", len(s)) + s for s in sps), bytes([len(pps)]),
               b"".join(struct.pack(">H", len(p)) + p for p in pps))
    avc1 = box(b"avc1", bytes(6), struct.pack(">H", 1), bytes(16), struct.pack(">HH", width, height),
               struct.pack(">II", 0x00480000, 0x00480000), bytes(4), struct.pack(">H", 1), bytes(32),
               struct.pack(">Hh", 24, -1), avcc)
    stsd = full(b"stsd", 0, 0, struct.pack(">I", 1), avc1)
    stts = full(b"stts", 0, 0, struct.pack(">III", 1, len(samples), 1))
    n_chunks = (len(samples) + chunk - 1) // chunk
    stsc_rows = [(1, chunk, 1)]
    if len(samples) % chunk:
        stsc_rows.append((n_chunks, len(samples) % chunk, 1))
        if n_chunks == 1:
            stsc_rows = stsc_rows[1:]
    stsc = full(b"stsc", 0, 0, struct.pack(">I", len(stsc_rows)), b"".join(struct.pack(">III", *r) for r in stsc_rows))
    stsz = full(b"stsz", 0, 0, struct.pack(">II", 0, len(samples)), b"".join(struct.pack(">I", len(s)) for s in samples))
    stss = full(b"stss", 0, 0, struct.pack(">II", 1, 1))

    def moov_with(offsets):
        if co64:
            stco = full(b"co64", 0, 0, struct.pack(">I", len(offsets)), b"".join(struct.pack(">Q", o) for o in offsets))
        else:
            stco = full(b"stco", 0, 0, struct.pack(">I", len(offsets)), b"".join(struct.pack(">I", o) for o in offsets))
        stbl = box(b"stbl", stsd, stts, stss, stsc, stsz, stco)
        dinf = box(b"dinf", full(b"dref", 0, 0, struct.pack(">I", 1), full(b"url ", 0, 1)))
        minf = box(b"minf", full(b"vmhd", 0, 1, bytes(8)), dinf, stbl)
        mdhd = full(b"mdhd", 0, 0, struct.pack(">IIIIHH", 0, 0, timescale, len(samples), 0x55C4, 0))
        hdlr = full(b"hdlr", 0, 0, bytes(4), b"vide", bytes(12), b"VideoHandler\x00")
        mdia = box(b"mdia", mdhd, hdlr, minf)
        matrix = struct.pack(">9I", 0x10000, 0, 0, 0, 0x10000, 0, 0, 0, 0x40000000)
        tkhd = full(b"tkhd", 0, 3, struct.pack(">IIIII", 0, 0, 1, 0, len(samples)), bytes(8), struct.pack(">HHHH", 0, 0, 0, 0),
                    matrix, struct.pack(">II", width << 16, height << 16))
        mvhd = full(b"mvhd", 0, 0, struct.pack(">IIII", 0, 0, timescale, len(samples)), struct.pack(">IH", 0x10000, 0x100),
                    bytes(10), matrix, bytes(24), struct.pack(">I", 2))
        return box(b"moov", mvhd, box(b"trak", tkhd, mdia))

    ftyp = box(b"ftyp", b"isom", struct.pack(">I", 512), b"isomiso2avc1mp41")
    moov = moov_with([0] * n_chunks)
    base = len(ftyp) + len(moov) + 8
    offsets, pos = [], base
    for c in range(n_chunks):
        offsets.append(pos)
        pos += sum(len(s) for s in samples[c * chunk:(c + 1) * chunk])
    moov = moov_with(offsets)
    return ftyp + moov + box(b"mdat", b"".join(samples))
