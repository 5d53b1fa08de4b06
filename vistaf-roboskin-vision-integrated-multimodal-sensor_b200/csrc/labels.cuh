// placeholder, filled in below
#pragma once
#include "common.cuh"
