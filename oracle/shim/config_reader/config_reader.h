// TEST INFRASTRUCTURE ONLY -- stand-in for ut-amrl/config-reader (needs Lua 5.1,
// absent from this image) so that /root/reference/src/util/kdtree.cc, which
// includes ../input/normal_computation.h, compiles UNCHANGED into oracle/_ref/.
// The CONFIG_* declarations of that header are never read on the KD-tree path;
// they become constants holding config/default_config.lua:144-156's values.
#ifndef NAUTILUS_ORACLE_SHIM_CONFIG_READER_
#define NAUTILUS_ORACLE_SHIM_CONFIG_READER_
#define CONFIG_DOUBLE(name, key) static const double CONFIG_##name = 0.0
#define CONFIG_FLOAT(name, key) static const float CONFIG_##name = 0.0f
#define CONFIG_INT(name, key) static const int CONFIG_##name = 0
#define CONFIG_BOOL(name, key) static const bool CONFIG_##name = false
#define CONFIG_STRING(name, key) static const char* const CONFIG_##name = ""
#endif
