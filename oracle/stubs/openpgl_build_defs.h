// Compile definitions the reference's cmake passes on the command line
// (src/ext/openpgl/openpgl/CMakeLists.txt:43-78), force-included by oracle/build_ref.sh.
#pragma once
#define OPENPGL_VSP_GUIDING
#define OPENPGL_IMAGE_SPACE_GUIDING_BUFFER
#define OPENPGL_DEVICE_TYPE_CPU_4
#define OPENPGL_BUILD
#define BUILD_SHARED
#ifndef OPENPGL_VERSION_MAJOR
#define OPENPGL_VERSION_MAJOR 0
#define OPENPGL_VERSION_MINOR 7
#define OPENPGL_VERSION_PATCH 0
#endif
#define OPENPGL_VERSION_STRING "0.7.0"
#define TBB_SUPPRESS_DEPRECATED_MESSAGES 1
