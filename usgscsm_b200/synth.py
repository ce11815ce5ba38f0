"""Synthetic model states for the hot-path workloads (no ISD / ALE involved).

A model state is what ``getModelState()`` of the reference writes: ``<MODEL_NAME>\\n{json}``
with the keys of ``UsgsAstroLsSensorModel::getModelMap`` (src/UsgsAstroLsSensorModel.cpp:345-474).
The orbit recipe follows jupyter_notebooks/OrbitalLsGeneration.ipynb of the reference
(circular orbit, nadir pointing), plus a small attitude wobble so that the 8-point Lagrange
interpolation of the quaternion table is exercised; LRO-NAC-like camera constants follow
tests/data/toughLroNacLineScan.json.

Every float array starts with a float literal: the reference types a JSON array by its
first element (src/Utilities.cpp:2760-2768).
"""
import json
import math

import numpy as np

RADIAL, TRANSVERSE, KAGUYALISM, DAWNFC, LROLROCNAC, CAHVOR, LUNARORBITER, RADTAN, KPLOSHADOWCAM = range(9)

MOON_RADIUS = 1737400.0


def _rot(axis, a):
    c, s = math.cos(a), math.sin(a)
    if axis == 0:
        return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)
    if axis == 1:
        return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def _quat_xyzw(m):
    """Rotation matrix -> unit quaternion (x, y, z, w), the storage order of m_quaternions."""
    t = m[0, 0] + m[1, 1] + m[2, 2]
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        w, x, y, z = 0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        s = math.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
        w, x, y, z = (m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s
    elif m[1, 1] > m[2, 2]:
        s = math.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
        w, x, y, z = (m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s
    else:
        s = math.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
        w, x, y, z = (m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s
    return np.array([x, y, z, w], dtype=np.float64)


def synthetic_dem(line, samp, amplitude=500.0):
    """h = A sin(2 pi l / 4096) cos(2 pi s / 1024) metres."""
    return amplitude * np.sin(2 * np.pi * line / 4096.0) * np.cos(2 * np.pi * samp / 1024.0)


def linescan_state(n_lines=52224, n_samples=5064, focal_length=699.62, pixel_pitch_inv=142.857,
                   altitude=50000.0, major_axis=MOON_RADIUS, minor_axis=MOON_RADIUS,
                   distortion_type=LROLROCNAC, dist_coeffs=(1.81e-5,), seed=20261104,
                   wobble=1e-5, wobble_hz=3.0, dt_ephem=1.0, dt_quat=0.05, line_time=None,
                   params=None, platform_flag=1, n_time_segments=1, gm=4.9028e12,
                   summing=1, starting_sample=0.0, derived=None, as_dict=False, polar=False):
    """A line-scan model state on a circular, tilted polar orbit with nadir pointing.
    `polar`: an exactly polar orbit instead (the orbit plane contains the body's axis; the
    ground track runs along a meridian, so a north-up DEM raster lines up with the image).

    `line_time` defaults to the value giving square ground pixels.  `n_time_segments` > 1
    splits the image into segments with slightly different line rates (exercises the
    upper_bound of getImageTime, LS:1543-1568).  `derived` = None leaves gsd/flyingHeight at
    their defaults so the model computes them itself (LS:283-285).
    """
    rng = np.random.default_rng(seed)
    r_orbit = major_axis + altitude
    v = math.sqrt(gm / r_orbit)
    omega = v / r_orbit
    ground_speed = omega * major_axis
    gsd_cross = altitude / (focal_length * pixel_pitch_inv) * summing
    if line_time is None:
        line_time = gsd_cross / ground_speed

    # line -> time segments (m_intTimeLines is 0.5-based as written by the reference's ISD path)
    seg_lines, seg_starts, seg_times = [], [], []
    t = 0.0
    per = n_lines // n_time_segments
    for k in range(n_time_segments):
        l0 = k * per
        rate = line_time * (1.0 + 0.01 * k)
        seg_lines.append(l0 + 0.5)
        seg_starts.append(t)
        seg_times.append(rate)
        nl = (n_lines - l0) if k == n_time_segments - 1 else per
        t += rate * nl
    total = t
    seg_starts = [s - total / 2.0 for s in seg_starts]  # time is relative to the image centre
    t_begin, t_end = -total / 2.0, total / 2.0

    tilt = _rot(2, 0.3 + 0.1 * rng.random()) @ _rot(0, 0.2 + 0.1 * rng.random())
    phase0 = 0.4 + rng.random()
    if polar:
        tilt = _rot(2, 0.3 + 0.1 * rng.random())
        phase0 = 0.2 + 0.6 * rng.random()   # latitude of the image centre, radians

    def frame(tt):
        th = phase0 + omega * tt
        radial = tilt @ np.array([math.cos(th), 0.0, math.sin(th)])
        intrack = tilt @ np.array([-math.sin(th), 0.0, math.cos(th)])
        cross = np.cross(radial, intrack)
        return radial, intrack, cross

    n_pos = int(math.ceil((t_end - t_begin) / dt_ephem)) + 1 + 8
    t0_ephem = t_begin - 4 * dt_ephem
    pos = np.empty((n_pos, 3))
    vel = np.empty((n_pos, 3))
    for i in range(n_pos):
        radial, intrack, _ = frame(t0_ephem + i * dt_ephem)
        pos[i] = r_orbit * radial
        vel[i] = v * intrack

    n_quat = int(math.ceil((t_end - t_begin) / dt_quat)) + 1 + 8
    t0_quat = t_begin - 4 * dt_quat
    ph = rng.random(3) * 2 * math.pi
    quat = np.empty((n_quat, 4))
    for i in range(n_quat):
        tt = t0_quat + i * dt_quat
        radial, intrack, cross = frame(tt)
        nominal = np.column_stack([intrack, cross, radial])  # camera x, y, z axes in body frame
        w = [wobble * math.sin(2 * math.pi * wobble_hz * (1 + 0.13 * k) * tt + ph[k]) for k in range(3)]
        q = _quat_xyzw(nominal @ _rot(0, w[0]) @ _rot(1, w[1]) @ _rot(2, w[2]))
        if i and float(np.dot(q, quat[i - 1])) < 0:
            q = -q
        quat[i] = q

    p = [0.0] * 16 if params is None else [float(x) for x in params]
    d = derived or {}
    state = {
        "m_modelName": "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL",
        "m_imageIdentifier": "synthetic_ls_%d" % seed,
        "m_sensorName": "SYNTHETIC_NAC",
        "m_nLines": int(n_lines),
        "m_nSamples": int(n_samples),
        "m_platformFlag": int(platform_flag),
        "m_intTimeLines": [float(x) for x in seg_lines],
        "m_intTimeStartTimes": [float(x) for x in seg_starts],
        "m_intTimes": [float(x) for x in seg_times],
        "m_startingEphemerisTime": float(t_begin),
        "m_centerEphemerisTime": 0.0,
        "m_detectorSampleSumming": int(summing),
        "m_detectorLineSumming": int(summing),
        "m_startingDetectorSample": float(starting_sample),
        "m_startingDetectorLine": 0.0,
        "m_ikCode": -85600,
        "m_focalLength": float(focal_length),
        "m_zDirection": 1.0,
        "m_distortionType": int(distortion_type),
        "m_opticalDistCoeffs": [float(x) for x in dist_coeffs],
        "m_iTransS": [0.0, 0.0, float(pixel_pitch_inv)],
        "m_iTransL": [0.0, float(pixel_pitch_inv), 0.0],
        "m_detectorSampleOrigin": float(n_samples * summing / 2.0 - 0.5),
        "m_detectorLineOrigin": 0.0,
        "m_majorAxis": float(major_axis),
        "m_minorAxis": float(minor_axis),
        "m_platformIdentifier": "SYNTH",
        "m_sensorIdentifier": "SYNTH_NAC",
        "m_minElevation": -1000.0,
        "m_maxElevation": 1000.0,
        "m_dtEphem": float(dt_ephem),
        "m_t0Ephem": float(t0_ephem),
        "m_dtQuat": float(dt_quat),
        "m_t0Quat": float(t0_quat),
        "m_numPositions": int(3 * n_pos),
        "m_numQuaternions": int(4 * n_quat),
        "m_referencePointXyz": [0.0, 0.0, 0.0],
        "m_gsd": float(d.get("gsd", 1.0)),
        "m_flyingHeight": float(d.get("flyingHeight", 1000.0)),
        "m_halfSwath": float(d.get("halfSwath", 1000.0)),
        "m_halfTime": float(d.get("halfTime", 10.0)),
        "m_positions": [float(x) for x in pos.reshape(-1)],
        "m_velocities": [float(x) for x in vel.reshape(-1)],
        "m_quaternions": [float(x) for x in quat.reshape(-1)],
        "m_currentParameterValue": p,
        "m_covariance": [0.0] * 256,
        "m_sunPosition": [1.5e11, 0.0, 0.0],
        "m_sunVelocity": [0.0, 0.0, 0.0],
    }
    if as_dict:
        return state
    return state_string(state)


def _orbit_tables(t_begin, t_end, dt_ephem, dt_quat, r_orbit, gm, rng, wobble, wobble_hz, pad=4):
    """Positions/velocities every dt_ephem and camera-to-body quaternions (x,y,z,w) every
    dt_quat over [t_begin, t_end] padded by `pad` nodes: circular tilted orbit, nadir pointing
    (camera x = in-track, y = cross-track, z = radial) plus a small attitude wobble."""
    v = math.sqrt(gm / r_orbit)
    omega = v / r_orbit
    tilt = _rot(2, 0.3 + 0.1 * rng.random()) @ _rot(0, 0.2 + 0.1 * rng.random())
    phase0 = 0.4 + rng.random()

    def frame(tt):
        th = phase0 + omega * tt
        radial = tilt @ np.array([math.cos(th), 0.0, math.sin(th)])
        intrack = tilt @ np.array([-math.sin(th), 0.0, math.cos(th)])
        return radial, intrack, np.cross(radial, intrack)

    n_pos = int(math.ceil((t_end - t_begin) / dt_ephem)) + 1 + 2 * pad
    t0_ephem = t_begin - pad * dt_ephem
    pos, vel = np.empty((n_pos, 3)), np.empty((n_pos, 3))
    for i in range(n_pos):
        radial, intrack, _ = frame(t0_ephem + i * dt_ephem)
        pos[i], vel[i] = r_orbit * radial, v * intrack
    n_quat = int(math.ceil((t_end - t_begin) / dt_quat)) + 1 + 2 * pad
    t0_quat = t_begin - pad * dt_quat
    ph = rng.random(3) * 2 * math.pi
    quat = np.empty((n_quat, 4))
    for i in range(n_quat):
        tt = t0_quat + i * dt_quat
        radial, intrack, cross = frame(tt)
        nominal = np.column_stack([intrack, cross, radial])
        w = [wobble * math.sin(2 * math.pi * wobble_hz * (1 + 0.13 * k) * tt + ph[k]) for k in range(3)]
        q = _quat_xyzw(nominal @ _rot(0, w[0]) @ _rot(1, w[1]) @ _rot(2, w[2]))
        if i and float(np.dot(q, quat[i - 1])) < 0:
            q = -q
        quat[i] = q
    return dict(pos=pos, vel=vel, quat=quat, t0_ephem=t0_ephem, t0_quat=t0_quat, omega=omega, v=v)


def pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096, overlap=2,
                    focal_length=6.0, pixel_pitch_inv=640.0, altitude=50000.0,
                    major_axis=MOON_RADIUS, minor_axis=MOON_RADIUS, distortion_type=RADIAL,
                    dist_coeffs=(-0.0113109, -1.44463e-4, -4.88754e-6), seed=3, wobble=1e-5,
                    wobble_hz=0.5, dt_ephem=1.0, dt_quat=0.25, exposure=0.039, params=None,
                    platform_flag=1, gm=4.9028e12, summing=1.0, flipped=False, reversed_order=False,
                    line_sign=-1.0, derived=None, as_dict=False):
    """A push-frame model state (keys of UsgsAstroPushFrameSensorModel::getModelMap, PF:332-466;
    shape of config C3 / tests/data/lroWacPushFrame.json): `n_framelets` framelets of
    `framelet_height` reduced lines each (the detector framelet has `overlap` more lines),
    one exposure per framelet, interframe delay chosen so consecutive framelets tile the ground."""
    rng = np.random.default_rng(seed)
    r_orbit = major_axis + altitude
    v = math.sqrt(gm / r_orbit)
    ground_speed = v / r_orbit * major_axis
    gsd = altitude / (focal_length * pixel_pitch_inv) * summing
    delay = framelet_height * gsd / ground_speed
    total = n_framelets * delay
    t_begin = -total / 2.0
    orb = _orbit_tables(t_begin, t_begin + total, dt_ephem, dt_quat, r_orbit, gm, rng, wobble, wobble_hz)
    p = [0.0] * 16 if params is None else [float(x) for x in params]
    d = derived or {}
    full_height = int(framelet_height * summing) + overlap
    state = {
        "m_modelName": "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL",
        "m_imageIdentifier": "synthetic_pf_%d" % seed,
        "m_sensorName": "SYNTHETIC_WAC",
        "m_nLines": int(n_framelets * (framelet_height + overlap)),
        "m_nSamples": int(n_samples),
        "m_platformFlag": int(platform_flag),
        "m_exposureDuration": float(exposure),
        "m_interframeDelay": float(delay),
        "m_frameletHeight": int(full_height),
        "m_frameletsFlipped": bool(flipped),
        "m_frameletsOrderReversed": bool(reversed_order),
        "m_startingEphemerisTime": float(t_begin),
        "m_centerEphemerisTime": 0.0,
        "m_detectorSampleSumming": float(summing),
        "m_detectorLineSumming": float(summing),
        "m_startingDetectorSample": 0.0,
        "m_startingDetectorLine": 0.0,
        "m_ikCode": -85621,
        "m_focalLength": float(focal_length),
        "m_zDirection": 1.0,
        "m_distortionType": int(distortion_type),
        "m_opticalDistCoeffs": [float(x) for x in dist_coeffs],
        "m_iTransS": [0.0, 0.0, float(pixel_pitch_inv)],
        "m_iTransL": [0.0, float(line_sign * pixel_pitch_inv), 0.0],
        "m_detectorSampleOrigin": float(n_samples * summing / 2.0),
        "m_detectorLineOrigin": float(full_height / 2.0),
        "m_majorAxis": float(major_axis),
        "m_minorAxis": float(minor_axis),
        "m_platformIdentifier": "SYNTH",
        "m_sensorIdentifier": "SYNTH_WAC",
        "m_minElevation": -1000.0,
        "m_maxElevation": 1000.0,
        "m_dtEphem": float(dt_ephem),
        "m_t0Ephem": float(orb["t0_ephem"]),
        "m_dtQuat": float(dt_quat),
        "m_t0Quat": float(orb["t0_quat"]),
        "m_numPositions": int(orb["pos"].size),
        "m_numQuaternions": int(orb["quat"].size),
        "m_referencePointXyz": [0.0, 0.0, 0.0],
        "m_gsd": float(d.get("gsd", 1.0)),
        "m_flyingHeight": float(d.get("flyingHeight", 1000.0)),
        "m_halfSwath": float(d.get("halfSwath", 1000.0)),
        "m_halfTime": float(d.get("halfTime", 10.0)),
        "m_positions": [float(x) for x in orb["pos"].reshape(-1)],
        "m_velocities": [float(x) for x in orb["vel"].reshape(-1)],
        "m_quaternions": [float(x) for x in orb["quat"].reshape(-1)],
        "m_currentParameterValue": p,
        "m_covariance": [0.0] * 256,
        "m_sunPosition": [1.5e11, 0.0, 0.0],
        "m_sunVelocity": [0.0, 0.0, 0.0],
        "m_numLinesOverlap": int(overlap),
        "m_reducedframeletHeight": int(framelet_height * summing),
        "m_nReducedLines": int(n_framelets * framelet_height),
    }
    if as_dict:
        return state
    return state_string(state)


def frame_state(n_lines=4096, n_samples=4096, focal_length=100.0, pixel_pitch_inv=100.0,
                altitude=50000.0, major_axis=MOON_RADIUS, minor_axis=MOON_RADIUS,
                distortion_type=RADIAL, dist_coeffs=(1e-4, -2e-7, 3e-10), seed=5, off_nadir=0.05,
                summing=1, starting_sample=0.0, starting_line=0.0, jitter=None, params_delta=None,
                unnormalised=1.0, as_dict=False):
    """A framing-camera model state (keys of UsgsAstroFrameSensorModel::getModelState,
    FR:1242-1288): one exposure from `altitude` above a random sub-spacecraft point, boresight
    (camera +z) tilted `off_nadir` rad from nadir.  `jitter` = (line_coeffs, sample_coeffs,
    line_time_step) adds the rolling-shutter polynomial of tests/data/jitterFramer-style states;
    `unnormalised` scales the stored quaternion (imageToGround normalises it, groundToImage
    does not: FR:2242-2286)."""
    rng = np.random.default_rng(seed)
    tilt = _rot(2, rng.random() * 6.0) @ _rot(0, rng.random() * 3.0)
    radial = tilt @ np.array([1.0, 0.0, 0.0])
    intrack = tilt @ np.array([0.0, 0.0, 1.0])
    zc = -radial
    yc = np.cross(zc, intrack)
    nominal = np.column_stack([intrack, yc, zc]) @ _rot(0, off_nadir) @ _rot(1, -0.7 * off_nadir) @ _rot(2, 0.3)
    q = _quat_xyzw(nominal) * unnormalised
    # geodetic-ish: put the camera above the ellipsoid along `radial`
    surf = 1.0 / math.sqrt((radial[0] ** 2 + radial[1] ** 2) / major_axis ** 2 + radial[2] ** 2 / minor_axis ** 2)
    pos = radial * (surf + altitude)
    p = [float(v) for v in pos] + [float(v) for v in q]
    if params_delta is not None:
        p = [a + float(b) for a, b in zip(p, params_delta)]
    state = {
        "m_modelName": "USGS_ASTRO_FRAME_SENSOR_MODEL",
        "m_majorAxis": float(major_axis),
        "m_minorAxis": float(minor_axis),
        "m_focalLength": float(focal_length),
        "m_startingDetectorSample": float(starting_sample),
        "m_startingDetectorLine": float(starting_line),
        "m_detectorSampleSumming": int(summing),
        "m_detectorLineSumming": int(summing),
        "m_focalLengthEpsilon": 1.0,
        "m_nLines": int(n_lines),
        "m_nSamples": int(n_samples),
        "m_ephemerisTime": 1000.0,
        "m_currentParameterValue": p,
        "m_ccdCenter": [float(n_lines * summing / 2.0), float(n_samples * summing / 2.0)],
        "m_spacecraftVelocity": [0.0, 1600.0, 0.0],
        "m_sunPosition": [1.5e11, 0.0, 0.0],
        "m_distortionType": int(distortion_type),
        "m_opticalDistCoeffs": [float(x) for x in dist_coeffs],
        "m_transX": [0.0, 1.0 / pixel_pitch_inv, 0.0],
        "m_transY": [0.0, 0.0, 1.0 / pixel_pitch_inv],
        "m_iTransS": [0.0, float(pixel_pitch_inv), 0.0],
        "m_iTransL": [0.0, 0.0, float(pixel_pitch_inv)],
        "m_imageIdentifier": "synthetic_frame_%d" % seed,
        "m_platformName": "SYNTH",
        "m_sensorName": "SYNTH_FRAMER",
        "m_collectionIdentifier": "",
        "m_currentParameterCovariance": [0.0] * 49,
        "m_minElevation": -1000.0,
        "m_maxElevation": 1000.0,
        "m_originalHalfLines": float(n_lines / 2.0),
        "m_originalHalfSamples": float(n_samples / 2.0),
        "m_pixelPitch": 1.0 / pixel_pitch_inv,
    }
    if jitter is not None:
        lj, sj, step = jitter
        state["m_lineJitter"] = [float(v) for v in lj]
        state["m_sampleJitter"] = [float(v) for v in sj]
        state["m_lineTimes"] = [float(step * (i - n_lines / 2.0)) + 1e-9 for i in range(n_lines)]
    if as_dict:
        return state
    return state_string(state)


def sar_state(n_lines=64576, n_samples=2367, exposure=0.004704421474, pixel_width=7.5,
              wavelength=0.12596322416750805, altitude=50000.0, near_slant=79942.38,
              major_axis=MOON_RADIUS, minor_axis=MOON_RADIUS, look="right", n_scale_sets=20,
              dt_ephem=0.25, seed=4, t_start=325441417.4304223, params=None, gm=4.9028e12,
              single_node=False, as_dict=False):
    """A SAR model state shaped like tests/data/orbitalSar.json after the reference's ISD load
    (keys of UsgsAstroSarSensorModel::getModelState, SAR:396-438): circular orbit, `n_scale_sets`
    cubic ground-range -> slant-range polynomials (slowly varying in time) fitted to the
    spherical geometry, absolute ephemeris times."""
    rng = np.random.default_rng(seed)
    r_orbit = major_axis + altitude
    total = n_lines * exposure
    t_end = t_start + total
    orb = _orbit_tables(t_start, t_end, dt_ephem, 10.0, r_orbit, gm, rng, 0.0, 1.0, pad=6)
    # cubic slant(ground range) for a sphere of radius major_axis seen from r_orbit
    R, r = major_axis, r_orbit
    th0 = math.acos((R * R + r * r - near_slant ** 2) / (2 * R * r))
    g = np.linspace(0.0, n_samples * pixel_width, 400)
    slant = np.sqrt(R * R + r * r - 2 * R * r * np.cos(th0 + g / R))
    base = np.polyfit(g, slant, 3)[::-1]
    times = t_start + total * np.arange(n_scale_sets) / max(n_scale_sets - 1, 1)
    coeffs = []
    for k in range(n_scale_sets):
        f = 1.0 + 2e-4 * math.sin(0.7 * k)
        coeffs += [float(base[0] * (1.0 + 1e-8 * k)), float(base[1] * f), float(base[2] / f), float(base[3])]
    pos, vel = orb["pos"], orb["vel"]
    if single_node:
        pos, vel = pos[len(pos) // 2: len(pos) // 2 + 1], vel[len(vel) // 2: len(vel) // 2 + 1]
    p = [0.0] * 6 if params is None else [float(x) for x in params]
    state = {
        "m_modelName": "USGS_ASTRO_SAR_SENSOR_MODEL",
        "m_imageIdentifier": "synthetic_sar_%d" % seed,
        "m_sensorName": "SYNTH_MINI_RF",
        "m_platformIdentifier": "SYNTH",
        "m_sensorIdentifier": "SYNTH_MINI_RF",
        "m_nLines": int(n_lines),
        "m_nSamples": int(n_samples),
        "m_exposureDuration": float(exposure),
        "m_scaledPixelWidth": float(pixel_width),
        "m_lookDirection": look,
        "m_wavelength": float(wavelength),
        "m_startingEphemerisTime": float(t_start),
        "m_centerEphemerisTime": float(t_start + total / 2.0),
        "m_endingEphemerisTime": float(t_end),
        "m_majorAxis": float(major_axis),
        "m_minorAxis": float(minor_axis),
        "m_minElevation": -1000.0,
        "m_maxElevation": 1000.0,
        "m_dtEphem": float(dt_ephem),
        "m_t0Ephem": float(orb["t0_ephem"]),
        "m_scaleConversionCoefficients": coeffs,
        "m_scaleConversionTimes": [float(t) for t in times],
        "m_positions": [float(x) for x in pos.reshape(-1)],
        "m_velocities": [float(x) for x in vel.reshape(-1)],
        "m_currentParameterValue": p,
        "m_referencePointXyz": [1.0, 0.0, 0.0],
        "m_covariance": [0.0] * 36,
        "m_sunPosition": [1.5e11, 0.0, 0.0],
        "m_sunVelocity": [0.0, 0.0, 0.0],
    }
    if as_dict:
        return state
    return state_string(state)


def state_string(state):
    """``<MODEL_NAME>\\n{json}`` as getModelState() writes it (LS:476-483)."""
    return state["m_modelName"] + "\n" + json.dumps(state)
