/* TEST INFRASTRUCTURE -- `recording`: a sampler plugin for the UNMODIFIED reference (built next to its own plugins by
   oracle/Makefile.ref) that behaves exactly like `independent` (src/samplers/independent.cpp:54-93: the same PCG32Sampler
   base, the same next_1d / next_2d) and additionally writes down, at every advance() -- which SamplingIntegrator::
   render_sample calls once per sample (src/librender/integrator.cpp:287) --

       seed value (m_base_seed + seed_offset: block_id * block_size^2 + morton index, integrator.cpp:198),
       index of the sample within that seed, PCG32 state after the sample, number of draws since the seed.

   That is the per-path information the north star asks to be bit-exact ("PCG32 states ... per path") and which the
   reference binary does not expose otherwise: with it the oracle and the CUDA path are compared with the REFERENCE's own
   run, not with each other (tests/test_reference_parity.py::test_per_path_states_against_the_reference).
   Records go to the file named by B2VP_RECORD_PATH when the process exits: 4 x uint64 per record. */
#include <mitsuba/core/properties.h>
#include <mitsuba/render/sampler.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <vector>

namespace {
struct Record { uint64_t seed, index, state, draws; };
struct Recorder {
    std::mutex mutex;
    std::vector<Record> records;
    ~Recorder() {
        const char *path = std::getenv("B2VP_RECORD_PATH");
        if (!path) return;
        if (FILE *f = std::fopen(path, "wb")) {
            std::fwrite(records.data(), sizeof(Record), records.size(), f);
            std::fclose(f);
        }
    }
    void add(const std::vector<Record> &r) { std::lock_guard<std::mutex> lock(mutex); records.insert(records.end(), r.begin(), r.end()); }
};
Recorder g_recorder;
}

NAMESPACE_BEGIN(mitsuba)

template <typename Float, typename Spectrum>
class RecordingSampler final : public PCG32Sampler<Float, Spectrum> {
public:
    MTS_IMPORT_BASE(PCG32Sampler, m_sample_count, m_base_seed, m_rng, seeded)
    MTS_IMPORT_TYPES()

    RecordingSampler(const Properties &props = Properties()) : Base(props) {
        if (!is_dynamic_array_v<Float>)
            Base::seed(PCG32_DEFAULT_STATE);
    }
    ~RecordingSampler() { flush(); }

    ref<Sampler<Float, Spectrum>> clone() override {
        RecordingSampler *sampler = new RecordingSampler();
        sampler->m_sample_count = m_sample_count;
        sampler->m_base_seed = m_base_seed;
        return sampler;
    }

    void seed(uint64_t seed_offset, size_t wavefront_size = 1) override {
        Base::seed(seed_offset, wavefront_size);
        m_seed_value = m_base_seed + seed_offset;
        m_index = 0;
        m_draws = 0;
    }

    Float next_1d(Mask active = true) override {
        if constexpr (!is_array_v<Float>) {
            if (active) ++m_draws;
        }
        return m_rng.template next_float<Float>(active);
    }

    Point2f next_2d(Mask active = true) override {
        Float f1 = next_1d(active), f2 = next_1d(active);
        return Point2f(f1, f2);
    }

    void advance() override {
        Base::advance();
        if constexpr (!is_array_v<Float>) {
            m_local.push_back({ m_seed_value, m_index++, (uint64_t) m_rng.state, m_draws });
            if (m_local.size() >= 4096) flush();
        }
    }

    std::string to_string() const override { return "RecordingSampler[]"; }
    MTS_DECLARE_CLASS()
private:
    void flush() { if (!m_local.empty()) { g_recorder.add(m_local); m_local.clear(); } }
    uint64_t m_seed_value = 0, m_index = 0, m_draws = 0;
    std::vector<Record> m_local;
};

MTS_IMPLEMENT_CLASS_VARIANT(RecordingSampler, Sampler)
MTS_EXPORT_PLUGIN(RecordingSampler, "Recording sampler (test infrastructure)");
NAMESPACE_END(mitsuba)
