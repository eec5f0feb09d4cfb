"""Minimal reader for the raw-appended VTU pieces pfc_vtu_write produces (tests only)."""
import re
import struct
import xml.etree.ElementTree as ET

import numpy as np

_DT = {"Float64": np.float64, "Int64": np.int64, "UInt8": np.uint8}


def read_piece(path):
    raw = open(path, "rb").read()
    cut = raw.index(b'<AppendedData encoding="raw">')
    start = raw.index(b"_", cut) + 1
    head = raw[:cut].decode() + "</VTKFile>"
    root = ET.fromstring(head)
    assert root.attrib["type"] == "UnstructuredGrid" and root.attrib["header_type"] == "UInt64"
    piece = root.find("UnstructuredGrid/Piece")
    out = {"n_points": int(piece.attrib["NumberOfPoints"]), "n_cells": int(piece.attrib["NumberOfCells"]), "point_data": {}}

    def arr(da):
        off = start + int(da.attrib["offset"])
        (nbytes,) = struct.unpack("<Q", raw[off:off + 8])
        a = np.frombuffer(raw[off + 8:off + 8 + nbytes], dtype=_DT[da.attrib["type"]])
        nc = int(da.attrib.get("NumberOfComponents", 1))
        return a.reshape(-1, nc) if nc > 1 else a

    out["points"] = arr(piece.find("Points/DataArray"))
    for da in piece.findall("Cells/DataArray"):
        out[da.attrib["Name"]] = arr(da)
    for da in piece.findall("PointData/DataArray"):
        out["point_data"][da.attrib["Name"]] = arr(da)
    tail = raw[cut:].decode(errors="ignore")
    assert tail.rstrip().endswith("</VTKFile>")
    return out


def read_record(path):
    root = ET.parse(path).getroot()
    assert root.attrib["type"] == "PUnstructuredGrid"
    g = root.find("PUnstructuredGrid")
    return {"pieces": [p.attrib["Source"] for p in g.findall("Piece")],
            "arrays": {a.attrib["Name"]: int(a.attrib.get("NumberOfComponents", 1)) for a in g.findall("PPointData/PDataArray")}}
