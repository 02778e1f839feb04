#define FT_GROUP 16
#include "host_run.inc"
