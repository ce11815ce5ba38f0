"""Model states in msgpack binary form (the reference's --output-binary-model-state,
bin/usgscsm_cam_test.cc:157-170, README.md:60-62): the engine's decoder against the
reference's own encoder (nlohmann::json::to_msgpack through oracle/_ref) and against the
python msgpack package, on CPU for the host logic and on the GPU for the models built."""
import ctypes as C
import json

import numpy as np
import pytest


def _states():
    from usgscsm_b200 import synth
    return {
        "linescan": synth.linescan_state(n_lines=512, n_samples=256, seed=5),
        "pushframe": synth.pushframe_state(n_framelets=16, framelet_height=16, n_samples=256),
        "frame": synth.frame_state(n_lines=256, n_samples=256),
        "sar": synth.sar_state(),
    }


def _json_part(state):
    return json.loads(state[state.index("{"):])


def _name(raw):
    from usgscsm_b200 import _capi
    buf = C.create_string_buffer(128)
    code = _capi.lib.usgscsm_b200_model_name_from_state(raw, len(raw), buf, 128)
    return code, buf.value.decode()


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_model_name_from_msgpack_state(ref, kind):
    import msgpack
    from oracle import ref as oracle_ref
    state = _states()[kind]
    expect = _json_part(state)["m_modelName"]
    packed_ref = oracle_ref.state_to_msgpack(state)          # nlohmann's encoder
    packed_py = msgpack.packb(_json_part(state), use_single_float=False)
    assert packed_ref[0] in (0xDE, 0xDF) or 0x80 <= packed_ref[0] <= 0x8F
    # the two encoders agree on the document (nlohmann may narrow exact floats to float32)
    assert msgpack.unpackb(packed_ref) == msgpack.unpackb(packed_py)
    assert _name(packed_ref) == (0, expect)
    assert _name(packed_py) == (0, expect)


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_writer_is_byte_identical_to_the_reference(ref, kind):
    """usgscsm_b200_state_to_msgpack against nlohmann::json::to_msgpack on the same state, and
    a round trip through the engine's own reader."""
    from oracle import ref as oracle_ref
    from usgscsm_b200 import plugin
    state = _states()[kind]
    ours = plugin.stateToMsgpack(state)
    assert ours == oracle_ref.state_to_msgpack(state)
    assert plugin.stateToMsgpack(ours) == ours            # msgpack in -> same msgpack out
    assert _name(ours) == (0, _json_part(state)["m_modelName"])


def test_writer_number_forms():
    import msgpack
    from usgscsm_b200 import plugin
    doc = {"m_modelName": "USGS_ASTRO_FRAME_SENSOR_MODEL",
           "ints": [0, 127, 128, 255, 256, 65535, 65536, 4294967295, 4294967296, -1, -32, -33,
                    -128, -129, -32768, -32769, -2147483648, -2147483649],
           "floats": [0.5, 0.1, 1e300, -2.5, 3.0e38, 1e39], "mixed": [1, 2.5, "x", None, True, [1, 2]],
           "s": "a" * 40, "long": "b" * 300, "empty": [], "obj": {"k": [1.0, 2]}}
    ours = plugin.stateToMsgpack(json.dumps(doc))
    assert msgpack.unpackb(ours) == doc
    # integer-valued float literals stay floats, integers stay integers
    back = msgpack.unpackb(ours)
    assert isinstance(back["obj"]["k"][0], float) and isinstance(back["obj"]["k"][1], int)


def test_malformed_msgpack_is_rejected():
    from usgscsm_b200 import _capi
    import msgpack
    good = msgpack.packb({"m_modelName": "USGS_ASTRO_FRAME_SENSOR_MODEL", "a": [1, 2.5, "x"]})
    assert _name(good)[0] == 0
    assert _name(good[:-3])[0] == _capi.ESTATE                 # truncated
    assert _name(good + b"\x00")[0] == _capi.ESTATE            # trailing bytes
    assert _name(msgpack.packb({"m_modelName": 7}))[0] == _capi.ESTATE
    assert _name(msgpack.packb({"x": 1}))[0] == _capi.ESTATE
    assert _name(bytes([0x81, 0xC4, 0x01, 0x00, 0x01]))[0] == _capi.ESTATE   # bin key


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_model_from_msgpack_equals_model_from_json(ref, kind):
    import usgscsm_b200 as ub
    from oracle import ref as oracle_ref
    state = _states()[kind]
    packed = oracle_ref.state_to_msgpack(state)
    plugin = ub.UsgsAstroPlugin()
    assert plugin.getModelNameFromModelState(packed) == _json_part(state)["m_modelName"]
    a = plugin.constructModelFromState(state)
    b = plugin.constructModelFromState(packed)
    assert a.getImageSize() == b.getImageSize()
    sz = a.getImageSize()
    rng = np.random.default_rng(2)
    line, samp = rng.uniform(0, sz.line, 5000), rng.uniform(0, sz.samp, 5000)
    ra = a.imageToGroundBatch(line, samp, 50.0)
    rb = b.imageToGroundBatch(line, samp, 50.0)
    for u, v in zip(ra, rb):
        assert np.array_equal(u, v, equal_nan=True)
    ga = a.groundToImageBatch(*ra[:3])
    gb = b.groundToImageBatch(*rb[:3])
    for u, v in zip(ga, gb):
        assert np.array_equal(u, v, equal_nan=True)


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_json_writer_round_trips_every_bit(kind):
    """usgscsm_b200_state_to_json (the text getModelState writes): JSON -> JSON keeps the
    document, msgpack -> JSON -> msgpack keeps the bytes, number kinds survive."""
    from usgscsm_b200 import plugin
    state = _states()[kind]
    text = plugin.stateToJson(state)
    head, body = text.split("\n", 1)
    assert head == _json_part(state)["m_modelName"]
    assert json.loads(body) == _json_part(state)
    packed = plugin.stateToMsgpack(state)
    assert plugin.stateToJson(packed) == text                    # either encoding in, same text out
    assert plugin.stateToMsgpack(text) == packed                 # and back to the same bytes


def test_json_writer_number_forms_and_escapes():
    from usgscsm_b200 import plugin
    doc = {"m_modelName": "USGS_ASTRO_FRAME_SENSOR_MODEL",
           "ints": [0, -1, 65536, 4294967296, -2147483649], "floats": [0.1, 1e300, -2.5, 5e-324, 1.0, -0.0],
           "mixed": [1, 2.5, "x", None, True, [1, 2.0]], "s": 'quote " backslash \\ newline \n tab \t',
           "empty": [], "obj": {"k": [1.0, 2]}}
    text = plugin.stateToJson(json.dumps(doc))
    back = json.loads(text.split("\n", 1)[1])
    assert back == doc
    assert isinstance(back["floats"][4], float) and isinstance(back["ints"][0], int)
    assert isinstance(back["obj"]["k"][0], float) and isinstance(back["obj"]["k"][1], int)
    assert str(back["floats"][5]) == "-0.0"
