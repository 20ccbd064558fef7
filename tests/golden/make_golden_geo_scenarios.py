"""Generate tests/golden/geo_scenarios.npz: the reference's own geodetic scenario groups
(/root/reference/geospatial/tests/geospatial_tests.py:67-215), captured by RUNNING the reference's test routine
`_run_documented_tests` with a recording wrapper around GeospatialConverter.geodetic_to_xyz / xyz_to_norm -- no scenario
is restated here.  Container only:
    python tests/golden/make_golden_geo_scenarios.py
Stored per scenario: the geodetic points exactly as the reference fed them (fp64 values of its fp32 tensors), the
reference's fp64 ECEF, its normalised coordinates and the bounding box they were normalised with.
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")

from geospatial.geo2xyz import GeospatialConverter  # noqa: E402  (reference)
import geospatial.tests.geospatial_tests as ref_tests  # noqa: E402  (reference)

records = []
orig_xyz, orig_norm = GeospatialConverter.geodetic_to_xyz, GeospatialConverter.xyz_to_norm


def rec_xyz(self, geo, *a, **k):
    out = orig_xyz(self, geo, *a, **k)
    records.append({"geo": geo.detach().cpu().double().numpy().copy(), "ecef": out[0].detach().cpu().numpy().copy()})
    return out


def rec_norm(self, xyz, *a, **k):
    out = orig_norm(self, xyz, *a, **k)
    r = records[-1]
    r["norm"] = out.detach().cpu().double().numpy().copy()
    r["bbox_min"] = self.bbox.min_point.detach().cpu().double().numpy().copy()
    r["bbox_span"] = self.bbox.span.detach().cpu().double().numpy().copy()
    return out


GeospatialConverter.geodetic_to_xyz, GeospatialConverter.xyz_to_norm = rec_xyz, rec_norm
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    ref_tests._run_documented_tests("cpu", torch.float64)
GeospatialConverter.geodetic_to_xyz, GeospatialConverter.xyz_to_norm = orig_xyz, orig_norm
names = [ln.split("|")[1].strip() for ln in buf.getvalue().splitlines() if ln.count("|") >= 9 and ln.split("|")[0].strip().isdigit()]
records = [r for r in records if "norm" in r]
assert len(records) == len(names) and len(records) >= 30, (len(records), len(names))
payload = {"names": np.array(names)}
for i, r in enumerate(records):
    for k, v in r.items():
        payload[f"s{i}_{k}"] = v
np.savez_compressed(os.path.join(HERE, "geo_scenarios.npz"), **payload)
print(len(records), "scenario groups,", sum(len(r["geo"]) for r in records), "points:", names[:3], "...")
