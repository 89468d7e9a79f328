// scene_build.h — host: flatten the ABI scene into the kernel parameter (scene_build.cpp).
#pragma once
#include <stddef.h>
#include <vector>
#include "scene_dev.h"
#include "../../include/gamer_b200.h"

// getWinding (galaxycomponent.h:156-165) of galaxy g as a certified table; see scene_build.cpp. Fills d.wt_s0 / wt_n and
// appends the coefficients to `coeffs` (d.wt_off = their offset); leaves wt_n = 0 when the table cannot be certified.
void gk_build_winding_table(const GamerGalaxy &g, DGalaxy &d, std::vector<double> &coeffs);
// the table evaluated as the device evaluates it (tests)
double gk_winding_from_table(const DGalaxy &d, const double *coeffs, float radius);

// returns GAMER_OK or GAMER_E_ARG (message in err). wtab_host (may be null) receives the winding tables of the galaxies;
// the caller copies them to the device and stores the device address in out->wtab.
int gk_build_scene(const GamerFrame *f, const GamerGalaxy *gal, const GamerComponent *comp, DScene *out, char *err,
                   size_t errlen, std::vector<double> *wtab_host = nullptr);
