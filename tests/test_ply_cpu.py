"""Host half of the mesh-file decoders, no GPU needed: mtsb_ply_probe / mtsb_serialized_probe parse the PLY header (and index a
mixed face element) / open the offset dictionary and inflate the mesh header — counts against the meshes the reference's loaders
produced (tests/golden/ply_meshes.npz), and the refusals."""
import numpy as np
import pytest

import ply_cases


def test_probes_agree_with_the_reference_loader(pkg, golden):
    B = pkg.binding
    gold = golden("ply_meshes.npz")
    for name, c in ply_cases.cases().items():
        assert B.ply_probe(c["image"]) == (gold[name + "_P"].shape[0], gold[name + "_T"].shape[0]), name
    for name, c in ply_cases.ser_cases(pkg.scenes).items():
        idx = c["kwargs"].get("shape_index", 0)
        assert B.serialized_probe(c["image"], idx) == (gold[name + "_P"].shape[0], gold[name + "_T"].shape[0]), name


def test_probes_refuse_what_the_decoder_cannot_take(pkg):
    B = pkg.binding
    img = ply_cases.cases()["tri_n_uv"]["image"]
    with pytest.raises(B.MtsbError, match="binary_little_endian"):
        B.ply_probe(img.replace(b"binary_little_endian", b"ascii"))
    with pytest.raises(B.MtsbError, match="neither all triangles nor all quads"):
        B.ply_probe(img[:-5])
    with pytest.raises(B.MtsbError, match="no 'ply' magic"):
        B.ply_probe(b"plx\n" + img[4:])
    with pytest.raises(B.MtsbError, match="no end_header"):
        B.ply_probe(img[:60])
    with pytest.raises(B.MtsbError, match="list property"):
        B.ply_probe(img.replace(b"property float u\n", b"property list uchar float u\n"))
    # a face with five vertices in a file whose size fits neither uniform layout
    hdr = b"ply\nformat binary_little_endian 1.0\nelement vertex 5\nproperty float x\nproperty float y\nproperty float z\nelement face 2\nproperty list uchar int vertex_indices\nend_header\n"
    verts = np.zeros((5, 3), "<f4").tobytes()
    tri = bytes([3]) + np.array([0, 1, 2], "<i4").tobytes()
    five = hdr.replace(b"element face 2", b"element face 3") + verts + tri + bytes([5]) + np.array([0, 1, 2, 3, 4], "<i4").tobytes() + tri
    with pytest.raises(B.MtsbError, match="has 5 vertices"):
        B.ply_probe(five)
    mixed = hdr + verts + bytes([3]) + np.array([0, 1, 2], "<i4").tobytes() + bytes([4]) + np.array([0, 1, 2, 3], "<i4").tobytes()
    assert B.ply_probe(mixed) == (5, 3)
    ser = ply_cases.ser_cases(pkg.scenes)["ser_v4_n_uv"]["image"]
    with pytest.raises(B.MtsbError, match="invalid file format"):
        B.serialized_probe(b"\x00\x00" + ser[2:])
    with pytest.raises(B.MtsbError, match="incompatible file version"):
        B.serialized_probe(ser[:2] + b"\x09\x00" + ser[4:])
    with pytest.raises(B.MtsbError, match="out of range"):
        B.serialized_probe(ser, 3)
    with pytest.raises(B.MtsbError, match="inflate|truncated"):
        B.serialized_probe(ser[:4] + b"\xff" * 64)
