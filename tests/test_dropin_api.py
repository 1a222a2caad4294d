"""The drop-in modules expose the reference's API surface: every class, public method, property and module-level
function of the hot-path modules recorded in tests/golden/api_manifest.json (dumped from the unmodified reference by
tests/golden/make_api_manifest.py) exists with the same parameter names, order and defaults.  CPU only."""
import importlib
import inspect
import json
import os

import pytest

from conftest import GOLDEN

MANIFEST = json.load(open(os.path.join(GOLDEN, "api_manifest.json")))
# module-private helpers of the reference that are not part of the surface scripts use
SKIP_FUNCTIONS = {"distortion": {"_process_soft_lim", "_process_rapp", "_process_toi"}}


@pytest.fixture(scope="module")
def mods():
    from m_mimo_ofdm_with_nonlinear_pa_sim_b200 import dropin
    dropin.install()
    return {name: importlib.import_module(name) for name in MANIFEST}


def params(fn):
    return [[p.name, None if p.default is inspect.Parameter.empty else repr(p.default)]
            for p in inspect.signature(fn).parameters.values()]


@pytest.mark.parametrize("module", sorted(MANIFEST))
def test_module_surface(mods, module):
    mod, want = mods[module], MANIFEST[module]
    assert os.path.basename(os.path.dirname(mod.__file__)) == "dropin"
    for fname, sig in want["functions"].items():
        if fname in SKIP_FUNCTIONS.get(module, ()):
            continue
        assert hasattr(mod, fname), "%s.%s missing" % (module, fname)
        assert params(getattr(mod, fname)) == sig, "%s.%s signature" % (module, fname)
    for cname, c in want["classes"].items():
        cls = getattr(mod, cname, None)
        assert inspect.isclass(cls), "%s.%s missing" % (module, cname)
        for base in c["bases"]:
            if base not in ("object", "ABC"):
                assert any(b.__name__ == base for b in cls.__mro__), "%s must derive from %s" % (cname, base)
        for mname, sig in c["methods"].items():
            meth = getattr(cls, mname, None)
            assert callable(meth), "%s.%s.%s missing" % (module, cname, mname)
            assert params(meth) == sig, "%s.%s.%s signature: %s vs %s" % (module, cname, mname, params(meth), sig)
        for prop in c["properties"]:
            assert isinstance(inspect.getattr_static(cls, prop), property), "%s.%s property" % (cname, prop)


def test_object_attributes_without_gpu(mods):
    """Constructors that need no numeric work behave like the reference's (attribute names scripts read)."""
    m = mods["modulation"].OfdmQamModem(constel_size=64, n_fft=4096, n_sub_carr=2048, cp_len=128)
    assert (m.n_bits_per_ofdm_sym, m.n_bits_per_symbol, m.constellation_size, m.constel_size) == (12288, 6, 64, 64)
    assert m.avg_symbol_power == 42.0 and m.avg_sample_power == 21.0 and m.alpha == 1 and m.precoding_mat is None
    assert list(m.constellation[:4]) == [-7 - 7j, -7 - 5j, -7 - 1j, -7 - 3j]
    with pytest.raises(ValueError):
        mods["modulation"].QamModem(32)
    with pytest.raises(ValueError):
        mods["modulation"].Modem([1, -1, 1j])
    d = mods["distortion"].SoftLimiter(3.0, 21.0)
    assert str(d) == "softlim" and abs(d.sat_pow - 21.0 * 10 ** 0.3) < 1e-12
    d.set_ibo(0.0)
    d.set_avg_sample_power(2.0)
    assert d.sat_pow == 2.0
    t = mods["distortion"].ThirdOrderNonLin(20.0, 21.0)
    assert abs(t.cubic_dist_coeff - 0.01 / 21.0) < 1e-15 and str(t) == "toi"
    tx = mods["transceiver"].Transceiver(modem=m, center_freq=int(3.5e9), carrier_spacing=int(15e3), impairment=d)
    assert tx.modem.alpha == m.calc_alpha(0.0) and (tx.cord_x, tx.tx_ant_gain_db) == (0, 0)
    arr = mods["antenna_array"].LinearArray(n_elements=8, base_transceiver=tx, center_freq=int(3.5e9), wav_len_spacing=0.5, cord_z=15)
    assert len(arr.array_elements) == 8 and arr.array_elements[0].modem is not tx.modem
    xs = [e.cord_x for e in arr.array_elements]
    assert abs(xs[0] + xs[-1]) < 1e-12 and abs((xs[1] - xs[0]) - 0.5 * 299792458.0 / 3.5e9) < 1e-12
    pl = mods["antenna_array"].PlanarRectangularArray(2, 3, tx, int(3.5e9), 0.5, cord_z=15)
    assert pl.n_elements == 6 and len(pl.array_elements) == 6
    nz = mods["noise"].Awgn(snr_db=10, seed=1234)
    assert nz.snr_db == 10 and nz.rng_gen.standard_normal() == __import__("numpy").random.default_rng(1234).standard_normal()
    assert [str(c()) for c in (mods["channel"].MisoLosFd, mods["channel"].MisoTwoPathFd)] == ["los", "two_path"]
    u = mods["utilities"]
    assert list(u.dec2bitarray([5, 3], 3)) == [1, 0, 1, 0, 1, 1] and u.bitarray2dec([1, 0, 1]) == 5
    assert abs(u.ebn0_to_snr(15.0, 2048, 2048, 64) - 22.781512503836435) < 1e-12
