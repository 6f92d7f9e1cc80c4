// aggregate.cu — neighbour lists + neighbour aggregation (placeholder; implemented in a later commit)
#include "common.cuh"
using namespace pigs;
extern "C" {
size_t pigs_neighbor_workspace_bytes(int64_t, int) { return 0; }
int pigs_neighbors_count(int64_t, int, const float*, const float*, int64_t*, void*, void*) { return set_error(-2, "not implemented", __FILE__, __LINE__); }
int pigs_neighbors_fill(int64_t, int, const float*, const float*, const float*, const int64_t*, int64_t*, float*, float*, float*, void*, void*) { return set_error(-2, "not implemented", __FILE__, __LINE__); }
int pigs_aggregate_forward(int64_t, int, int, int, int, const float*, const float*, const float*, const float*, const float*, const float*, const int64_t*, const int64_t*, const float*, const float*, const float*, float*, float*, float*, float*, void*) { return set_error(-2, "not implemented", __FILE__, __LINE__); }
size_t pigs_aggregate_backward_workspace_bytes(int64_t, int, int, int) { return 0; }
int pigs_aggregate_backward(int64_t, int, int, int, int, int64_t, const float*, const float*, const float*, const float*, const float*, const float*, const int64_t*, const int64_t*, const float*, const float*, const float*, const float*, const float*, const float*, const float*, float*, float*, float*, float*, float*, float*, void*, void*) { return set_error(-2, "not implemented", __FILE__, __LINE__); }
}
