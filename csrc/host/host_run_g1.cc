#define FT_GROUP 1
#include "host_run.inc"
