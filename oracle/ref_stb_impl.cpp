// TEST INFRASTRUCTURE: one translation unit that instantiates the stb single-header libraries the
// reference's ImageLoader / HDR writer link against (the reference does this inside its own
// visualizer TUs, which are GL-only and not part of the headless oracle build).
#define STB_IMAGE_WRITE_IMPLEMENTATION
#include "stb_image_write.h"
