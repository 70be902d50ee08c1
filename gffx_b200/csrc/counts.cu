// counts.cu — depth / coverage pipelines (K6/K7).
#include "engine_impl.cuh"

namespace gffx {

struct CountsState {
    int kind = 0;
    double ms = 0;
    uint32_t launches = 0;
};

void counts_release(EngineImpl* E) { delete E->counts; E->counts = nullptr; }
void counts_begin_impl(EngineImpl* E, int kind) { throw Err(E_STATE, "depth/coverage engine not built yet"); }
void counts_add_impl(EngineImpl* E, const uint32_t*, const uint32_t*, const uint32_t*, uint64_t, uint32_t) {
    throw Err(E_STATE, "depth/coverage engine not built yet");
}
void counts_finish_impl(EngineImpl* E, CountRows*) { throw Err(E_STATE, "depth/coverage engine not built yet"); }
double counts_ms_impl(const EngineImpl* E) { return E->counts ? E->counts->ms : 0.0; }
uint32_t counts_launches_impl(const EngineImpl* E) { return E->counts ? E->counts->launches : 0; }

}  // namespace gffx
