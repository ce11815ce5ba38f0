// PROJ link stubs for the oracle build (TEST INFRASTRUCTURE -- not part of the product).
// Only /root/reference/src/UsgsAstroProjectedSensorModel.cpp (out of scope, SURVEY.md §2
// row 12) calls PROJ; the library itself needs sqlite3 and cannot be built here.  The six
// entry points it names are given bodies so that the plugin links; creating a projection
// fails (returns null), which makes the projected model unconstructible -- as intended.
#include <proj.h>

extern "C" {
PJ_CONTEXT *proj_context_create(void) { return nullptr; }
PJ_CONTEXT *proj_context_destroy(PJ_CONTEXT *) { return nullptr; }
PJ *proj_create(PJ_CONTEXT *, const char *) { return nullptr; }
PJ *proj_create_crs_to_crs_from_pj(PJ_CONTEXT *, const PJ *, const PJ *, PJ_AREA *,
                                   const char *const *) { return nullptr; }
PJ *proj_destroy(PJ *) { return nullptr; }
PJ_COORD proj_trans(PJ *, PJ_DIRECTION, PJ_COORD c) { return c; }
}
