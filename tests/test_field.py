"""Dense-field evaluation (SPECIFICATIONS.MD:1963-2088): the patch-ordered inference kernel
against the general path over the same np.linspace grid -- coordinates and features must be bit-identical."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

RES = [2 ** i for i in range(4, 20)]


def florida_converter(de, dev):
    from _synth import FLORIDA_BBOX_GEO as B, FLORIDA_T_RANGE
    conv = de.GeospatialConverter(device=str(dev), t_range=FLORIDA_T_RANGE)
    lats = torch.linspace(B["south"], B["north"], 5, dtype=torch.float64, device=dev)
    lons = torch.linspace(B["west"], B["east"], 5, dtype=torch.float64, device=dev)
    alts = torch.tensor([B["alt_min"], B["alt_max"]], dtype=torch.float64, device=dev)
    conv.update_bbox(torch.cartesian_prod(lats, lons, alts))
    return conv


def linspace_points(lat_range, lon_range, n_lat, n_lon, alt, dev):
    """The reference's query construction: for lat in np.linspace(...): for lon in np.linspace(...)."""
    lats, lons = np.linspace(*lat_range, n_lat), np.linspace(*lon_range, n_lon)
    geo = np.stack([np.repeat(lats, n_lon), np.tile(lons, n_lat), np.full(n_lat * n_lon, alt)], 1)
    return torch.from_numpy(geo).to(dev)


@pytest.mark.parametrize("n_lat,n_lon,table_dtype,T", [(37, 53, "float16", 2 ** 16), (128, 96, "float32", 2 ** 14),
                                                        (1, 1, "float16", 2 ** 14), (16, 1, "float32", 2 ** 14),
                                                        (300, 257, "float16", 2 ** 19)])
def test_field_features_bit_identical_to_general_path(n_lat, n_lon, table_dtype, T):
    import deepearth_b200 as de
    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    L = 16 if T == 2 ** 19 else 8
    cfg = de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=T, n_spatial_levels=L,
                             n_temporal_levels=L, spatial_resolutions=RES[:L], temporal_resolutions=RES[:L],
                             grid4d_mode="decomposed", grid4d_table_dtype=table_dtype)
    enc = de.Grid4DEncoder(cfg).to(dev).eval()
    conv = florida_converter(de, dev)
    lat_range, lon_range, alt, when = (28.1, 28.9), (-81.9, -80.95), 12.5, 1.5e9
    feats, xyzt = enc.encode_field(conv, lat_range, lon_range, n_lat, n_lon, altitude=alt, time=when, return_coords=True)
    geo = linspace_points(lat_range, lon_range, n_lat, n_lon, alt, dev)
    t = torch.full((n_lat * n_lon,), when, dtype=torch.float64, device=dev)
    want_xyzt = conv.geodetic_to_xyzt_norm(geo, t)
    assert torch.equal(xyzt, want_xyzt), "generated coordinates differ from the prologue kernel on np.linspace points"
    assert torch.equal(conv.field_xyzt(lat_range, lon_range, n_lat, n_lon, alt, when), want_xyzt)
    with torch.no_grad():
        want = enc.encode_features(want_xyzt)
    assert feats.shape == want.shape
    assert torch.equal(feats, want), float((feats - want).abs().max())
    emb = enc.forward_field(conv, lat_range, lon_range, n_lat, n_lon, altitude=alt, time=when)
    assert emb.shape == (n_lat, n_lon, 32) and bool(torch.isfinite(emb).all())


def test_field_reference_geometry_uses_generated_coordinates():
    """Mode R (nearest, xyz + t encoders): coordinates generated on the device, general forward."""
    import deepearth_b200 as de
    dev = torch.device("cuda:0")
    torch.manual_seed(4)
    enc = de.Grid4DEncoder(de.DeepEarthConfig(hidden_dim=32, n_heads=4, n_layers=1, hash_table_size=2 ** 14)).to(dev).eval()
    conv = florida_converter(de, dev)
    feats = enc.encode_field(conv, (28.2, 28.8), (-81.8, -81.0), 40, 25, altitude=3.0, time=1.4e9)
    geo = linspace_points((28.2, 28.8), (-81.8, -81.0), 40, 25, 3.0, dev)
    t = torch.full((1000,), 1.4e9, dtype=torch.float64, device=dev)
    with torch.no_grad():
        want = enc.encode_features(conv.geodetic_to_xyzt_norm(geo, t))
    assert torch.equal(feats, want)


def test_timestamps_from_date_columns_match_python_datetime():
    """SPECIFICATIONS.MD:740-768: datetime(y, m, d, h, mi, int(s)).timestamp() (+ microseconds), evaluated for UTC."""
    import calendar
    import deepearth_b200 as de
    from datetime import datetime, timezone
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(11)
    n = 4000
    y = rng.integers(1901, 2200, n)
    m = rng.integers(1, 13, n)
    d = np.array([rng.integers(1, calendar.monthrange(int(a), int(b))[1] + 1) for a, b in zip(y, m)])
    h, mi = rng.integers(0, 24, n), rng.integers(0, 60, n)
    s = np.round(rng.uniform(0, 59.999999, n), 6)
    y[:4], m[:4], d[:4] = [2000, 1900 + 4, 2024, 1970], [2, 2, 2, 1], [29, 29, 29, 1]          # leap days, the epoch
    h[3] = mi[3] = 0
    s[3] = 0.0
    comp = np.stack([y, m, d, h, mi, s], 1).astype(np.float64)
    want = []
    for row in comp:
        dt = datetime(int(row[0]), int(row[1]), int(row[2]), int(row[3]), int(row[4]), int(row[5]), tzinfo=timezone.utc)
        whole = calendar.timegm(dt.timetuple())
        want.append(whole + int((row[5] % 1) * 1e6) / 1e6)
    want = np.array(want)
    conv = de.GeospatialConverter(device=str(dev), t_range=(1262304000.0, 1767139200.0))
    got = conv.timestamps_from_components(torch.from_numpy(comp).to(dev))
    assert np.array_equal(got.cpu().numpy(), want)
    tn = conv.timestamps_from_components(torch.from_numpy(comp).to(dev), normalise=True)
    want_n = ((want - 1262304000.0) / (1767139200.0 - 1262304000.0)).astype(np.float32)
    assert np.array_equal(tn.cpu().numpy(), want_n)
    # integer columns with defaults: year only -> June 15th, 12:00 (config.default_month / day / hour)
    yo = torch.tensor([[2015], [1999]], dtype=torch.int32, device=dev)
    got2 = conv.timestamps_from_components(yo).cpu().numpy()
    assert list(got2) == [calendar.timegm((2015, 6, 15, 12, 0, 0)), calendar.timegm((1999, 6, 15, 12, 0, 0))]
