// host_graph.h - the graph.pkl emitter shared by the host front end (host_graph.cu) and the device front end
// (graph_device.cu): both hand it nodes in first-appearance order and adjacency lists in edge-insertion order.
#pragma once
#include <stdint.h>

#define RCQ_GRAPH_FIRST_VISIT 0xFFFFFFFFu

int rcq_graph_emit(const char* path, int32_t n_samples, const char* const* sample_names, int64_t n_nodes,
                   const int32_t* node_sample, const int32_t* node_gene, const int64_t* start, const int32_t* nbr,
                   const uint32_t* erel, int64_t n_edges, int64_t counts[3]);
