"""ISD -> model state on the host (usgscsm_b200_state_from_isd; no CUDA device needed).

The product converter (usgscsm_b200/csrc/isd_to_state.cpp) against the oracle = the reference's
own constructStateFromIsd running on the restated ALE subset: same keys, same types, numbers
equal to rounding (the two sides order the Hermite / SLERP arithmetic differently, so the last
bits may differ: tolerance 1e-11 relative, stated here).  Then the reference's own asserted
values that pin the conversion itself (tests/SarTests.cpp:126-138), and the failure behaviour."""
import ctypes as C
import json

import numpy as np
import pytest

import ref_goldens as G

REL_TOL = 1e-11


def product_state(name=None, text=None, model_name=None, image_id="img"):
    from usgscsm_b200 import _capi
    L = _capi.lib
    raw = (G.isd_text(name) if text is None else text).encode()
    ident = (name or image_id).encode()
    need = C.c_size_t()
    mn = None if model_name is None else model_name.encode()
    rc = L.usgscsm_b200_state_from_isd(raw, len(raw), mn, ident, None, 0, C.byref(need))
    if rc != 0:
        raise RuntimeError("%d: %s" % (rc, L.usgscsm_b200_last_error().decode()))
    buf = C.create_string_buffer(need.value)
    assert L.usgscsm_b200_state_from_isd(raw, len(raw), mn, ident, buf, need.value, C.byref(need)) == 0
    s = buf.raw[:need.value].decode()
    head, body = s.split("\n", 1)
    return head, json.loads(body)


@pytest.mark.parametrize("name", G.ISD_FIXTURES)
def test_state_matches_reference_converter(ref, name):
    isd = json.loads(G.isd_text(name))
    isd["image_identifier"] = name
    want = json.loads(ref.state_from_isd(json.dumps(isd)))
    head, got = product_state(name)
    assert head == want["m_modelName"] == got["m_modelName"]
    # m_parameterType (a list of strings) never reaches the reference's VariantMap (U:2747-2775)
    assert set(got) - {"m_parameterType"} == set(want)
    for key, w in want.items():
        g = got[key]
        if isinstance(w, (str, bool)):
            assert g == w, key
            continue
        assert isinstance(g, list) == isinstance(w, list), key
        wa, ga = np.atleast_1d(np.array(w, dtype=float)), np.atleast_1d(np.array(g, dtype=float))
        assert wa.shape == ga.shape, key
        if wa.size == 0:
            continue
        # the reference's VariantMap types a value by its JSON literal: int must stay int
        first_w = w[0] if isinstance(w, list) else w
        first_g = g[0] if isinstance(g, list) else g
        assert isinstance(first_w, int) == isinstance(first_g, int), key
        nan = np.isnan(wa)
        assert np.array_equal(nan, np.isnan(ga)), key
        scale = max(np.abs(wa[~nan]).max(initial=0.0), 1e-300)
        assert np.abs(wa - ga)[~nan].max(initial=0.0) <= REL_TOL * scale, key


def test_sar_state_goldens_product():
    """tests/SarTests.cpp:126-138 through the PRODUCT's converter (see test_oracle_isd_golden)."""
    _, st = product_state("orbitalSar")
    t0, dt = st["m_t0Ephem"], st["m_dtEphem"]
    f = (G.SAR_STATE["time"] - t0) / dt
    for key, gold in (("m_positions", "position"), ("m_velocities", "velocity")):
        tab = np.array(st[key]).reshape(-1, 3)
        got = tab[0] + (tab[1] - tab[0]) * f
        exp, rel = G.SAR_STATE[gold]
        assert np.all(np.abs(got - np.array(exp)) <= rel * np.abs(exp)), (key, got, exp)


def test_resampling_shapes():
    """LS:2883-2895: >= 25 uniformly spaced ephemeris nodes; LS:3046-3053: quaternions on the
    merged attitude grid (NAC: 1773 pointing nodes + 2 body nodes, the ends coincide)."""
    _, st = product_state("constVelocityLineScan")
    assert st["m_numPositions"] == 75 and len(st["m_velocities"]) == 75
    assert st["m_numQuaternions"] == 36
    assert st["m_dtEphem"] == pytest.approx(1.6 / 24) and st["m_t0Ephem"] == pytest.approx(-0.8)
    assert st["m_majorAxis"] == 10.0 and st["m_gsd"] == 1.0 and st["m_flyingHeight"] == 1000.0
    _, st = product_state("toughLroNacLineScan")
    assert st["m_numPositions"] == 3 * 173 and st["m_distortionType"] == 4
    assert st["m_opticalDistCoeffs"] == [1.81e-05]
    _, st = product_state("orbitalPushFrame")
    assert st["m_reducedframeletHeight"] == st["m_frameletHeight"] - st["m_numLinesOverlap"]
    assert st["m_nReducedLines"] == (st["m_nLines"] // st["m_frameletHeight"]) * st["m_reducedframeletHeight"]
    _, st = product_state("jitterFramer")
    assert len(st["m_currentParameterValue"]) == 7 and len(st["m_lineTimes"]) == 16


@pytest.mark.parametrize("name", ["empty", "hello"])
def test_unusable_isd_is_refused(name):
    """tests/PluginTests.cpp: an empty object / a JSON string is not constructible."""
    with pytest.raises(RuntimeError) as e:
        product_state(name)
    assert e.value.args[0].startswith("-2")  # USGSCSM_B200_ESTATE


def test_missing_key_is_named():
    isd = json.loads(G.isd_text("orbitalLineScan"))
    del isd["body_rotation"]
    with pytest.raises(RuntimeError) as e:
        product_state(text=json.dumps(isd))
    assert "body rotation" in e.value.args[0]
    isd = json.loads(G.isd_text("orbitalLineScan"))
    with pytest.raises(RuntimeError):  # a frame converter refuses a line-scan ISD's missing keys? no:
        product_state(text=json.dumps(isd), model_name="USGS_ASTRO_SAR_SENSOR_MODEL")
    with pytest.raises(RuntimeError):
        product_state(text=json.dumps(isd), model_name="NOT_A_MODEL")


def test_python_mirror_constructStateFromIsd():
    import usgscsm_b200 as ub
    s = ub.UsgsAstroPlugin().constructStateFromIsd(G.isd_image("orbitalFramer"))
    head, body = s.split("\n", 1)
    assert head == "USGS_ASTRO_FRAME_SENSOR_MODEL"
    assert json.loads(body)["m_imageIdentifier"] == "orbitalFramer"
