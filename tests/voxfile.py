"""TEST HELPER — writes MagicaVoxel .vox files (Ephtracy's published format) so that the importer can be exercised on
files we own: SIZE / XYZI per model, optional RGBA palette, optional MATL dictionaries, plus chunks a loader must skip."""
import struct


def _chunk(cid: bytes, content: bytes, children: bytes = b"") -> bytes:
    return cid + struct.pack("<ii", len(content), len(children)) + content + children


def _dict(d) -> bytes:
    out = struct.pack("<i", len(d))
    for k, v in d.items():
        kb, vb = k.encode(), str(v).encode()
        out += struct.pack("<i", len(kb)) + kb + struct.pack("<i", len(vb)) + vb
    return out


def write_vox(models, palette=None, matl=None, version=150, extra_chunks=True) -> bytes:
    """models: list of (size3, voxels (n,4) uint8 [x, y, z, colourIndex]); palette: 255 ABGR words for indices 1..255 or None;
    matl: {paletteIndex: {"_type": "_metal", "_rough": 0.3, ...}}."""
    body = b""
    if extra_chunks:
        body += _chunk(b"PACK", struct.pack("<i", len(models)))
    for size, vox in models:
        body += _chunk(b"SIZE", struct.pack("<iii", *size))
        body += _chunk(b"XYZI", struct.pack("<i", len(vox)) + bytes(bytearray(int(v) for row in vox for v in row)))
    if extra_chunks:
        body += _chunk(b"nTRN", _dict({"_name": "root"}) + struct.pack("<iiii", 1, -1, 0, 1) + _dict({}))
    if palette is not None:
        assert len(palette) == 255
        body += _chunk(b"RGBA", b"".join(struct.pack("<I", int(c)) for c in palette) + struct.pack("<I", 0))
    for idx, props in (matl or {}).items():
        body += _chunk(b"MATL", struct.pack("<i", idx) + _dict(props))
    return b"VOX " + struct.pack("<i", version) + _chunk(b"MAIN", b"", body)
