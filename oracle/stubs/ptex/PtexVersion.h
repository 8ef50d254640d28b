// Test-infrastructure stub (NOT product code): what the reference's cmake generates from
// src/ext/ptex/src/ptex/PtexVersion.h.in with PTEX_VER "v2.4.X" (src/ext/ptex/CMakeLists.txt:60-74).
#ifndef PtexVersion_h
#define PtexVersion_h
#define PtexAPIVersion 4
#define PtexFileMajorVersion 1
#define PtexFileMinorVersion 4
#define PtexLibraryMajorVersion 2
#define PtexLibraryMinorVersion 4
#define PTEX_NAMESPACE Ptex
#define make_version_ns_(major, minor) v##major##_##minor
#define make_version_ns(major, minor) make_version_ns_(major, minor)
#define PTEX_VERSION_NAMESPACE make_version_ns(PtexLibraryMajorVersion, PtexLibraryMinorVersion)
#define PTEX_NAMESPACE_BEGIN        \
    namespace PTEX_NAMESPACE {      \
    namespace PTEX_VERSION_NAMESPACE {} \
    using namespace PTEX_VERSION_NAMESPACE; \
    namespace PTEX_VERSION_NAMESPACE {
#define PTEX_NAMESPACE_END }}
#endif
