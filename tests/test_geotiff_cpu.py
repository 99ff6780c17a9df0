"""SURVEY.md 8(f) next-4, on-disk format: the (Geo)TIFF reader against files
written by an independent encoder (libtiff through Pillow: LZW, DEFLATE,
PackBits, predictors 2 / 3, multiple strips) and the writer's tiled / DEFLATE
layouts against an independent decoder (Pillow again).  The reference reads
and writes through rasterio / GDAL (ocsmesh/raster.py:502-531,
ocsmesh/utils.py:1932-1971)."""
import numpy as np
import pytest

from ocsmesh_b200 import geotiff

PIL = pytest.importorskip("PIL.Image")
from PIL import features  # noqa: E402

libtiff = pytest.mark.skipif(not features.check("libtiff"),
                             reason="Pillow without libtiff")


def _dem(h, w, dtype=np.float32, seed=0):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    z = 30 * np.sin(x / 37.0) * np.cos(y / 23.0) + 0.01 * x + rng.normal(0, 0.05, (h, w))
    return z.astype(dtype)


@libtiff
@pytest.mark.parametrize("compression,predictor", [
    ("tiff_lzw", 1), ("tiff_lzw", 2), ("tiff_lzw", 3),
    ("tiff_adobe_deflate", 1), ("tiff_adobe_deflate", 3), ("tiff_deflate", 1),
    ("packbits", 1), (None, 1),
])
def test_reads_libtiff_float(tmp_path, compression, predictor):
    a = _dem(301, 517)
    p = tmp_path / "a.tif"
    info = {317: predictor} if predictor != 1 else {}
    PIL.fromarray(a, mode="F").save(p, compression=compression, tiffinfo=info)
    meta, bands = geotiff.read_geotiff(p)
    assert (meta.height, meta.width, meta.count) == (301, 517, 1)
    assert bands.dtype == np.float32
    np.testing.assert_array_equal(np.asarray(bands[0]), a)
    # the file really is what the case says (not silently stored uncompressed)
    with PIL.open(p) as im:
        want = {None: (1,), "tiff_lzw": (5,), "tiff_adobe_deflate": (8,),
                "tiff_deflate": (8, 32946), "packbits": (32773,)}[compression]
        assert im.tag_v2[259] in want
        assert im.tag_v2.get(317, 1) == predictor


@libtiff
@pytest.mark.parametrize("mode,dtype", [("I;16", np.uint16), ("L", np.uint8), ("I", np.int32)])
def test_reads_libtiff_integers_lzw_predictor2(tmp_path, mode, dtype):
    a = (_dem(200, 333, np.float64, seed=3) * 20 + 900).astype(dtype)
    p = tmp_path / "i.tif"
    PIL.fromarray(a, mode=mode).save(p, compression="tiff_lzw", tiffinfo={317: 2})
    meta, bands = geotiff.read_geotiff(p)
    assert bands.dtype == np.dtype(dtype)
    np.testing.assert_array_equal(np.asarray(bands[0]), a)


@libtiff
def test_lzw_table_reset_on_noise(tmp_path):
    # incompressible data fills the 4096-entry table many times (Clear codes)
    a = np.random.default_rng(7).integers(0, 256, (600, 700), dtype=np.uint8)
    p = tmp_path / "n.tif"
    PIL.fromarray(a, mode="L").save(p, compression="tiff_lzw")
    np.testing.assert_array_equal(np.asarray(geotiff.read_geotiff(p)[1][0]), a)


@pytest.mark.parametrize("tile,compress,predictor", [
    (64, None, 1), (128, "deflate", 1), (64, "deflate", 3), (None, "deflate", 2),
    (None, "deflate", 3), (256, "deflate", 2),
])
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_writer_layouts_round_trip_and_pillow(tmp_path, tile, compress, predictor, dtype):
    a = _dem(203, 331, dtype, seed=5)
    tr = (10.0, 0.0, 500000.0, 0.0, -10.0, 4102030.0)
    p = tmp_path / "w.tif"
    geotiff.write_geotiff(p, a, tr, epsg=32618, nodata=-9999.0, tile=tile,
                          compress=compress, predictor=predictor)
    meta, bands = geotiff.read_geotiff(p)
    assert meta.transform == tr and meta.epsg == 32618 and meta.nodata == -9999.0
    assert bands.dtype == np.dtype(dtype)
    np.testing.assert_array_equal(np.asarray(bands[0]), a)
    if dtype == np.float32 and features.check("libtiff"):
        # an independent decoder agrees (Pillow reads float32 as mode F)
        with PIL.open(p) as im:
            np.testing.assert_array_equal(np.array(im), a)


def test_writer_multiband_tiled(tmp_path):
    a = np.stack([_dem(100, 150, seed=k) for k in range(3)])
    p = tmp_path / "m.tif"
    geotiff.write_geotiff(p, a, (1.0, 0, 0.0, 0, -1.0, 100.0), tile=32, compress="deflate",
                          predictor=3)
    meta, bands = geotiff.read_geotiff(p)
    assert meta.count == 3
    np.testing.assert_array_equal(np.asarray(bands), a)


def test_raster_opens_compressed_tiled_file(tmp_path):
    import ocsmesh_b200 as ob
    a = _dem(180, 240, seed=9)
    tr = (10.0, 0.0, 500000.0, 0.0, -10.0, 4101800.0)
    p = tmp_path / "r.tif"
    geotiff.write_geotiff(p, a, tr, epsg=32618, tile=64, compress="deflate", predictor=3)
    r = ob.Raster(p, chunk_size=100, overlap=2)
    np.testing.assert_array_equal(r.get_values(band=1), a)
    win = list(r.iter_windows())[1]
    np.testing.assert_array_equal(
        r.get_values(window=win, band=1),
        a[win.row_off:win.row_off + win.height, win.col_off:win.col_off + win.width])


def test_errors(tmp_path):
    p = tmp_path / "bad.tif"
    a = _dem(40, 50)
    with pytest.raises(ValueError, match="predictor"):
        geotiff.write_geotiff(p, a.astype(np.int16), (1, 0, 0, 0, -1, 40), compress="deflate",
                              predictor=3)
    with pytest.raises(ValueError, match="multiple of 16"):
        geotiff.write_geotiff(p, a, (1, 0, 0, 0, -1, 40), tile=50)
    from ocsmesh_b200 import _lib
    with pytest.raises(_lib.OcsbError, match="corrupt"):
        _lib.tiff_decode(5, bytes([0xff, 0xff, 0xff, 0xff, 0xff]), 100)
    with pytest.raises(_lib.OcsbError, match="more than"):
        # PackBits: twenty runs of 128 x 'A' do not fit 10 bytes
        _lib.tiff_decode(32773, bytes([0x81, 0x41] * 20), 10)
