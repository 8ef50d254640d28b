// Hand-written equivalent of the cmake-configured openpgl/defines.h for the
// reference's build options (src/ext/CMakeLists.txt:262-273): VSP guiding ON,
// image-space guiding buffer ON, radiance caches OFF, compression OFF.
#pragma once
#define OPENPGL_VSP_GUIDING
#define OPENPGL_IMAGE_SPACE_GUIDING_BUFFER
