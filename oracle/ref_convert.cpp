// oracle/ref_convert.cpp -- TEST INFRASTRUCTURE ONLY.
// Runs the reference converter (tools/convert_icon/convert_icon.cpp, included where it lies, its own main() untouched) on
// stand-in NetCDF files (oracle/stub_nc/netcdf.h) with `.ic` output switched on.  The converter has no command-line switch
// for that (convertToIC{false}, convert_icon.cpp:22; flags parsed :121-161), so a static initialiser behind its globals
// sets the flag -- and maxLayers from REF_CONVERT_MAX_LAYERS -- before main() runs.
// Built by `make ref` into oracle/_ref/convert_icon_ref; usage: the converter's own arguments.
#include <climits>
#include <cstdlib>

#include "convert_icon.cpp"

namespace {
struct SwitchOnIcOutput {
  SwitchOnIcOutput()
  {
    g_appState.convertToIC = true;
    g_appState.convertToUMesh = false;
    if (const char *m = std::getenv("REF_CONVERT_MAX_LAYERS")) g_appState.maxLayers = std::atoi(m);
  }
} switchOnIcOutput;
}
