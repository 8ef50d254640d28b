// Hand-written equivalent of the cmake-configured openpgl/version.h
// (src/ext/openpgl/CMakeLists.txt:25: project version 0.7.0).
#pragma once
#define OPENPGL_VERSION_MAJOR 0
#define OPENPGL_VERSION_MINOR 7
#define OPENPGL_VERSION_PATCH 0
#define OPENPGL_VERSION "0.7.0"
