#include "internal.cuh"
