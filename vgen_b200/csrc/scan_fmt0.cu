#define VG_FMT_INST 0
#include "scan_fmt.inc"
