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
unsigned long long az_forest_uid(az_forest* f);           // unique per created forest, never reused (unlike its address)
void az_net_forget_forest(unsigned long long forest_uid); // az_net.cu: drop the step graphs captured for a destroyed forest
// The forest's own noise buffer f64 [n_trees][A] (device); az_forest_use_device_noise(f, 1) makes the searches read it
// (the self-play engine fills it on the device), 0 turns noise off.
double* az_forest_noise_dev(az_forest* f);
int az_forest_use_device_noise(az_forest* f, int on);
// Evaluator identity: cached answers belong to ONE evaluator; a different tag clears the evaluation cache.
int az_forest_set_eval_tag(az_forest* f, unsigned long long tag, cudaStream_t s);
int az_forest_check_eval_tag(az_forest* f, unsigned long long tag);
int az_search_burst(const AzForestDev* d, int iters, int steps_done);   // launches to queue before polling the done counter
struct az_selfplay;
int az_selfplay_service(az_selfplay* sp, cudaStream_t s);        // az_selfplay.cu: finished slots move and re-arm (capturable)
az_forest* az_selfplay_forest(az_selfplay* sp);
void az_selfplay_count_plies(az_selfplay* sp, long long n);
int az_forest_cache_tick(az_forest* f, cudaStream_t s);          // one unit of cache age without a synchronised begin_move (continuous plies)
unsigned long long az_hash_eval_tag(uint64_t seed, int qbits);   // evaluator identity of the hash evaluator (az_eval_hash)
