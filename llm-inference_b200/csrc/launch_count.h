// Process-wide counter of kernel launches issued by this library (b200_launch_count()).
// bench.py reports it as `gpu_launches`; graph replays add the node count of the graph.
#pragma once
#include <stdint.h>
namespace b200 {
void count_launch(uint64_t n = 1);
}
