// ALE shim (see Vectors.h).  Only the enumerator NAMES are used by the reference
// (src/Utilities.cpp:1865-1895); the numeric order matters only on the ISD path.
#ifndef ORACLE_SHIM_ALE_DISTORTION_H
#define ORACLE_SHIM_ALE_DISTORTION_H
namespace ale {
enum DistortionType {
  TRANSVERSE, RADIAL, KAGUYALISM, DAWNFC, LROLROCNAC, CAHVOR, LUNARORBITER, RADTAN,
  KPLOSHADOWCAM
};
}
#endif
