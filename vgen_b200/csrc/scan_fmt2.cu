#define VG_FMT_INST 2
#include "scan_fmt.inc"
