#define VG_FMT_INST 4
#include "scan_fmt.inc"
