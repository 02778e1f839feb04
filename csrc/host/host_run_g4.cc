#define FT_GROUP 4
#include "host_run.inc"
