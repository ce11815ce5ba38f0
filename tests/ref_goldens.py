"""The values the REFERENCE'S OWN TESTS assert for models built from its ISD fixtures
(/root/reference/tests/*.cpp; fixtures copied to tests/golden/ref_data/).  Restated here so that
the same checks run (a) against the oracle = reference code + restated ALE
(tests/test_oracle_isd_golden.py, CPU) and (b) against the CUDA engine through the product's own
ISD -> state converter (tests/test_isd_golden_gpu.py).

Each entry: (fixture, operation, inputs, expected, tolerance, citation).
  i2g : imageToGround((line, samp), height)      -> (x, y, z)
  g2i : groundToImage((x, y, z), prec)           -> (line, samp)
  rt  : imageToGround then groundToImage         -> the pixel again, within tolerance [px];
        inputs = (line, samp, height, prec)
"""
import os

DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_data")


def isd_image(name):
    """Path of the (absent) image file whose sidecar is the ISD, as tests/Fixtures.h passes it."""
    return os.path.join(DATA, name + ".img")


def isd_text(name):
    with open(os.path.join(DATA, name + ".json")) as f:
        return f.read()


ISD_FIXTURES = ["constVelocityLineScan", "orbitalLineScan", "flippedOrbitalLineScan",
                "toughLroNacLineScan", "ctxLineScan", "orbitalPushFrame", "lroWacPushFrame",
                "orbitalFramer", "simpleFramerISD", "jitterFramer", "dawnfc_radtan", "orbitalSar"]

I2G = [
    # tests/LineScanCameraTests.cpp:49-55 Center, :70-76 OffBody
    ("constVelocityLineScan", (8.0, 8.0), 0.0, (9.99999500000, 0.0, 0.00999999500), 1e-10, "LineScanCameraTests.cpp:49-55"),
    ("constVelocityLineScan", (0.0, 4.0), 0.0, (0.04799688020, -7.99961602495, 16.00004799688), 1e-10, "LineScanCameraTests.cpp:70-76"),
    # :215-221 OrbitalLineScanSensorModel.Center
    ("orbitalLineScan", (8.5, 8.0), 0.0, (999999.67004351015, 0.0, -812.35021438796923), 1e-9, "LineScanCameraTests.cpp:215-221"),
    # tests/FrameCameraTests.cpp:35-91
    ("simpleFramerISD", (7.5, 7.5), 0.0, (10.0, 0.0, 0.0), 1e-8, "FrameCameraTests.cpp:35-41"),
    ("simpleFramerISD", (7.5, 6.5), 0.0, (9.80194018, 0.0, 1.98039612), 1e-8, "FrameCameraTests.cpp:43-49"),
    ("simpleFramerISD", (15.0, 0.0), 0.0, (0.44979759, -14.99325304, 14.99325304), 1e-8, "FrameCameraTests.cpp:52-58"),
    ("simpleFramerISD", (0.0, 15.0), 0.0, (0.44979759, 14.99325304, -14.99325304), 1e-8, "FrameCameraTests.cpp:59-65"),
    ("simpleFramerISD", (0.0, 0.0), 0.0, (0.44979759, 14.99325304, 14.99325304), 1e-8, "FrameCameraTests.cpp:67-73"),
    ("simpleFramerISD", (15.0, 15.0), 0.0, (0.44979759, -14.99325304, -14.99325304), 1e-8, "FrameCameraTests.cpp:84-90"),
    ("orbitalFramer", (8.0, 8.0), 0.0, (1000000.0, 0.0, 0.0), 1e-8, "FrameCameraTests.cpp:109-115"),
    # tests/PushFrameCameraTests.cpp:94-99
    ("orbitalPushFrame", (6.0, 8.0), 0.0, (1000000.0, 0.0, 0.0), 1e-9, "PushFrameCameraTests.cpp:94-99"),
    # tests/SarTests.cpp:105-111
    ("orbitalSar", (500.0, 500.0), 0.0, (-1520427.2924246688, -450240.98096817418, 710030.04690435971), 1e-3, "SarTests.cpp:105-111"),
]

G2I = [
    # tests/PushFrameCameraTests.cpp:101-106
    ("orbitalPushFrame", (1000000.0, 0.0, 0.0), 0.001, (6.0, 8.0), 0.001, "PushFrameCameraTests.cpp:101-106"),
    # tests/SarTests.cpp:113-118
    ("orbitalSar", (-1520427.2924246688, -450240.98096817418, 710030.04690435971), 0.001, (500.0, 500.0), 1e-3, "SarTests.cpp:113-118"),
]

ROUND_TRIP = (
    # tests/LineScanCameraTests.cpp:58-68 Inversion
    [("constVelocityLineScan", (8.5, 8.0, 0.0, 0.001), 1e-3, "LineScanCameraTests.cpp:58-68")] +
    # :223-237 Inversion, :248-259 InversionHeight, :261-272 InversionReallyHigh (all 16 lines)
    [("orbitalLineScan", (l + 0.5, 8.0, 0.0, 0.001), 0.1, "LineScanCameraTests.cpp:223-237") for l in range(16)] +
    [("orbitalLineScan", (l + 0.5, 8.0, 100.0, 0.001), 0.1, "LineScanCameraTests.cpp:248-259") for l in range(16)] +
    [("orbitalLineScan", (l + 0.5, 8.0, 4900.0, 0.001), 0.002, "LineScanCameraTests.cpp:261-272") for l in range(16)] +
    # :279-285 OppositeFlightDetector
    [("flippedOrbitalLineScan", (8.7, 8.0, 0.0, 0.001), 0.001, "LineScanCameraTests.cpp:279-285")] +
    # tests/FrameCameraTests.cpp:97-103 Inversion (EXPECT_DOUBLE_EQ: 4 ulp)
    [("simpleFramerISD", (9.0, 9.0, 0.0, 0.001), 4e-15, "FrameCameraTests.cpp:97-103")] +
    # :209-219 JitterFrameSensorModel.Inversion: every integer pixel 0..16 x 0..16, height 0.001
    [("jitterFramer", (float(l), float(s), 0.001, 0.001), 0.001, "FrameCameraTests.cpp:209-219")
     for l in range(17) for s in range(17)]
)

# tests/PushFrameCameraTests.cpp:109-125 Inversion: all 360 x 16 pixel centres, 10 px tolerance
PUSHFRAME_INVERSION = ("orbitalPushFrame", 360, 16, 10.0, "PushFrameCameraTests.cpp:109-125")

# tests/SarTests.cpp:126-149: Lagrange (order 2, far outside the table) of m_positions /
# m_velocities / range coefficients at time -0.0025; pins the ISD -> state conversion itself
SAR_STATE = dict(
    time=-0.0025,
    position=((212048065766.25, 64616398201.375, 488643469101.125), 1e-10),
    velocity=((-427443125.77587891, -144589932.36868286, 203454745.77563477), 1e-10),
    citation="SarTests.cpp:126-138",
)

# tests/LineScanCameraTests.cpp:178-183 getSunPositionLagrange: sun position at t = -0.6
SUN_LS = ("orbitalLineScan", -0.6, (125.0, 125.0, 125.0), 1e-11, "LineScanCameraTests.cpp:178-183")
