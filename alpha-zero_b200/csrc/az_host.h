// Host-side helpers shared by the translation units of libaz_b200.so.
#pragma once
#include <algorithm>
#include <string>
#include "az_forest.h"

int az_fail(int code, const std::string& msg);
int az_make_game(int game, int rows, int cols, AzGame* out);
uint64_t az_tuple2_hash(uint64_t r, uint64_t c);

struct az_forest;
// Accessors used by the net / self-play units (the struct itself lives in az_forest.cu).
AzForestDev* az_forest_dev(az_forest* f);
