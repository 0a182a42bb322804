#define VG_FMT_INST 5
#include "scan_fmt.inc"
