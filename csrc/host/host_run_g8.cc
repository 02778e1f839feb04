#define FT_GROUP 8
#include "host_run.inc"
