// Native-engine extras reachable through the C ABI (cudapoa_b200_get_stats / cudapoa_b200_max_poas).
#pragma once

#include "cudapoa_b200.h"

#include <cstdint>

namespace claraparabricks
{
namespace genomeworks
{
namespace cudapoa
{

class EngineStatsSource
{
public:
    virtual ~EngineStatsSource()                              = default;
    virtual void fill_stats(cudapoa_b200_stats* out) const    = 0;
    virtual int32_t max_poas() const                          = 0;
    // Benchmark hook: re-run the device work of the last generate_poa() `iterations` times on inputs that are
    // already resident in device memory; returns the CUDA-event time of all iterations in ms (< 0: nothing staged).
    virtual double relaunch(int32_t iterations) = 0;
    // Same device work once more, with a CUDA event between the launches: milliseconds spent in the fill launches and
    // in the graph launches of the phase-split schedule (a batch on the single launch reports everything as fill_ms).
    // Returns the number of launches timed, < 0 if nothing has been staged.
    virtual int32_t relaunch_phases(double* fill_ms, double* graph_ms) = 0;
    // relaunch() and relaunch_phases() in one: `iterations` passes with events between the launches; returns the
    // CUDA-event time of all of them, the sums per kernel in *fill_ms / *graph_ms, the launch count in *launches.
    virtual double relaunch_timed(int32_t iterations, double* fill_ms, double* graph_ms, int32_t* launches) = 0;
    // Per-phase totals of the last relaunch_timed(): fill, traceback, fusion, topological sort, consensus / MSA (ms).
    virtual int32_t last_phase_ms(double* out, int32_t cap) const = 0;
    // Scheduling hint: this batch shares its device with other batches in flight (several streams).  A lone batch of
    // less than about two waves of windows is served faster by the single launch; overlapping batches by the
    // phase-split schedule whatever their size.
    virtual void set_overlapped(bool overlapped) = 0;
    // Development aid: copy a window's device state to the host.  region 0 = its arena slice (graph, records, scratch),
    // 1 = its score area (move codes, spilled rows).  Returns the region's size in bytes; copies min(size, cap).
    virtual int64_t debug_read_window(int32_t window, int32_t region, void* out, int64_t cap) = 0;
};

} // namespace cudapoa
} // namespace genomeworks
} // namespace claraparabricks
