// scene_build.h — host: flatten the ABI scene into the kernel parameter (scene_build.cpp).
#pragma once
#include <stddef.h>
#include "scene_dev.h"
#include "../../include/gamer_b200.h"

// returns GAMER_OK or GAMER_E_ARG (message in err)
int gk_build_scene(const GamerFrame *f, const GamerGalaxy *gal, const GamerComponent *comp, DScene *out, char *err,
                   size_t errlen);
