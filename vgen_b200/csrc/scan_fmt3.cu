#define VG_FMT_INST 3
#include "scan_fmt.inc"
