#define VG_FMT_INST 1
#include "scan_fmt.inc"
