#define FT_GROUP 2
#include "host_run.inc"
