// rowsplit.cu — placeholder (filled in below)
#include "common.cuh"
