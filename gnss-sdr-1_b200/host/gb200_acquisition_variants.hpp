// Host mirrors of the sibling PCPS blocks that run the same Doppler loop as pcps_acquisition, on top of
// the C ABI:
//   B200_Pcps_Tong_Acquisition        pcps_tong_acquisition_cc         (src/algorithms/acquisition/gnuradio_blocks/pcps_tong_acquisition_cc.cc)
//   B200_Galileo_Pcps_8ms_Acquisition galileo_pcps_8ms_acquisition_cc  (.../galileo_pcps_8ms_acquisition_cc.cc)
//   B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf  galileo_e5a_noncoherentIQ_acquisition_caf_cc (.../galileo_e5a_noncoherent_iq_acquisition_caf_cc.cc)
//   B200_Pcps_Acquisition_Fine_Doppler  pcps_acquisition_fine_doppler_cc (.../pcps_acquisition_fine_doppler_cc.cc)
// The blocks use fft_size = sampled_ms * samples_per_ms with NO zero padding, count their Doppler bins inclusively
// ("doppler <= doppler_max") and build the wipe-off tables with their own phase-step expression, so the bins are handed to
// the engine as an explicit list (gb200_acq_set_doppler_grid).  The device does the per-bin arithmetic; the counters,
// thresholds and Gnss_Synchro fields stay here.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <complex>
#include <cstdint>
#include <vector>

#include "../../include/gnss_b200.h"
#include "gb200_acquisition.hpp"

namespace gb200
{
struct AcqVariantConf
{
    uint32_t sampled_ms = 1;
    uint32_t doppler_max = 5000;
    uint32_t doppler_step = 250;
    int64_t fs_in = 4000000;
    int samples_per_ms = 4000;
    int samples_per_code = 4000;
    float threshold = 0.0f;
    uint32_t max_dwells = 1;                                          // 8 ms block
    uint32_t tong_init_val = 1, tong_max_val = 2, tong_max_dwells = 20;  // Tong block
    bool bit_transition_flag = false, both_signal_components = false;    // E5a block
    int caf_window_hz = 0, zero_padding = 0;                             // E5a block
    int item_type = GB200_ITEM_32FC;
};

namespace detail
{
inline gb200_acq_cfg variant_cfg(const AcqVariantConf& c, bool keep_grid_dwells)
{
    gb200_acq_cfg g{};
    g.sampled_ms = c.sampled_ms;
    g.ms_per_code = c.sampled_ms;  // fft_size == consumed samples: no zero padding in these blocks
    g.samples_per_chip = 1;
    g.max_dwells = keep_grid_dwells ? 0x7fffffffu : 1u;  // Tong: the engine never ends the attempt by itself
    g.doppler_max = c.doppler_max;
    g.doppler_step = c.doppler_step;
    g.fs_in = c.fs_in;
    g.samples_per_ms = static_cast<float>(c.samples_per_ms);
    g.samples_per_code = static_cast<float>(c.samples_per_code);
    g.use_CFAR_algorithm_flag = 1;
    g.item_type = c.item_type;
    g.threshold = c.threshold;
    return g;
}
inline uint32_t inclusive_bins(uint32_t doppler_max, uint32_t doppler_step)
{
    uint32_t n = 0;
    for (auto d = -static_cast<int32_t>(doppler_max); d <= static_cast<int32_t>(doppler_max); d += static_cast<int32_t>(doppler_step)) n++;
    return n;
}
}  // namespace detail

// ------------------------------------------------------------------------------------------------ Tong detector
class B200_Pcps_Tong_Acquisition
{
public:
    B200_Pcps_Tong_Acquisition(gb200_service* s, const AcqVariantConf& c, int stream_id) : s_(s), c_(c), stream_(stream_id)
    {
        const gb200_acq_cfg g = detail::variant_cfg(c, true);
        ok_ = gb200_acq_create(s_, &g, &h_) == GB200_OK;
        if (ok_) init();
    }
    ~B200_Pcps_Tong_Acquisition()
    {
        if (h_) gb200_acq_destroy(h_);
    }
    B200_Pcps_Tong_Acquisition(const B200_Pcps_Tong_Acquisition&) = delete;
    B200_Pcps_Tong_Acquisition& operator=(const B200_Pcps_Tong_Acquisition&) = delete;
    bool ok() const { return ok_; }
    void set_gnss_synchro(AcqSynchro* p) { synchro_ = p; }
    void set_threshold(float t) { c_.threshold = t; }
    void set_doppler_max(uint32_t v)
    {
        c_.doppler_max = v;
        init();
    }
    void set_doppler_step(uint32_t v)
    {
        c_.doppler_step = v;
        init();
    }
    // = init() (pcps_tong_acquisition_cc.cc:183-215): phase_step_rad = GPS_TWO_PI * doppler / float(fs) evaluated in double, stored as float
    void init()
    {
        const double GPS_TWO_PI = 6.283185307179586;
        bins_ = detail::inclusive_bins(c_.doppler_max, c_.doppler_step);
        std::vector<float> steps(bins_);
        for (uint32_t i = 0; i < bins_; i++)
            {
                const int32_t doppler = -static_cast<int32_t>(c_.doppler_max) + static_cast<int32_t>(c_.doppler_step * i);
                const float phase_step_rad = GPS_TWO_PI * doppler / static_cast<float>(c_.fs_in);
                steps[i] = phase_step_rad;
            }
        ok_ = ok_ && gb200_acq_set_doppler_grid(h_, bins_, steps.data()) == GB200_OK && gb200_acq_set_scaled_accumulation(h_, 1) == GB200_OK;
    }
    bool set_local_code(const std::complex<float>* code, size_t n_items)
    {
        return gb200_acq_set_local_code(h_, reinterpret_cast<const float*>(code), n_items) == GB200_OK;
    }
    // = set_state(1) (:226-252): counters and grid cleared
    void set_state(int state)
    {
        state_ = state;
        if (state == 1)
            {
                dwell_count_ = 0;
                tong_count_ = static_cast<int32_t>(c_.tong_init_val);
                mag_ = 0.0f;
                gb200_acq_reset(h_);
            }
    }
    void reset() { set_state(1); }
    uint32_t fft_size() const { return gb200_acq_fft_size(h_); }
    uint32_t num_doppler_bins() const { return bins_; }
    int32_t tong_count() const { return tong_count_; }
    uint32_t dwell_count() const { return dwell_count_; }
    float test_statistics() const { return test_statistics_; }
    float input_power() const { return input_power_; }
    uint32_t code_phase_index() const { return indext_; }

    // one pass through general_work state 1 (:290-395) on the fft_size items starting at samp_count.
    // Returns the event: 1 positive, 2 negative, 0 keep dwelling (call again with the next block), -1 error.
    int run(uint64_t samp_count)
    {
        if (state_ != 1) return 0;
        gb200_acq_result r;
        if (gb200_acq_run(h_, stream_, samp_count, &r) != GB200_OK) return -1;
        dwell_count_++;
        mag_ = r.peak_mag;  // d_mag: maximum of the accumulated, normalised grid
        input_power_ = r.input_power;
        indext_ = r.code_phase_index;
        if (synchro_)
            {
                synchro_->Acq_delay_samples = static_cast<double>(r.code_phase_index % static_cast<uint32_t>(c_.samples_per_code));
                synchro_->Acq_doppler_hz = static_cast<double>(-static_cast<int32_t>(c_.doppler_max) + static_cast<int32_t>(c_.doppler_step * r.doppler_bin));
                synchro_->Acq_samplestamp_samples = r.acq_samplestamp;
                synchro_->Acq_doppler_step = c_.doppler_step;
            }
        test_statistics_ = mag_;
        int state = 1;
        if (test_statistics_ > c_.threshold * dwell_count_)
            {
                tong_count_++;
                if (tong_count_ == static_cast<int32_t>(c_.tong_max_val)) state = 2;
            }
        else
            {
                tong_count_--;
                if (tong_count_ == 0) state = 3;
            }
        if (dwell_count_ >= c_.tong_max_dwells) state = 3;
        if (state == 1) return 0;
        state_ = 0;  // d_active = false; d_state = 0
        if (synchro_) synchro_->Flag_valid_acquisition = (state == 2);
        return state == 2 ? 1 : 2;
    }

private:
    gb200_service* s_;
    AcqVariantConf c_;
    gb200_acq* h_ = nullptr;
    AcqSynchro* synchro_ = nullptr;
    int stream_;
    bool ok_ = false;
    int state_ = 0;
    uint32_t bins_ = 0, dwell_count_ = 0, indext_ = 0;
    int32_t tong_count_ = 0;
    float mag_ = 0.0f, input_power_ = 0.0f, test_statistics_ = 0.0f;
};

// ------------------------------------------------------------------------------------------------ Galileo E1 8 ms (codes A / B)
class B200_Galileo_Pcps_8ms_Acquisition
{
public:
    B200_Galileo_Pcps_8ms_Acquisition(gb200_service* s, const AcqVariantConf& c, int stream_id) : s_(s), c_(c), stream_(stream_id)
    {
        const gb200_acq_cfg g = detail::variant_cfg(c, false);
        ok_ = gb200_acq_create(s_, &g, &h_[0]) == GB200_OK && gb200_acq_create(s_, &g, &h_[1]) == GB200_OK;
        if (ok_) init();
    }
    ~B200_Galileo_Pcps_8ms_Acquisition()
    {
        for (auto* h : h_)
            if (h) gb200_acq_destroy(h);
    }
    B200_Galileo_Pcps_8ms_Acquisition(const B200_Galileo_Pcps_8ms_Acquisition&) = delete;
    B200_Galileo_Pcps_8ms_Acquisition& operator=(const B200_Galileo_Pcps_8ms_Acquisition&) = delete;
    bool ok() const { return ok_; }
    void set_gnss_synchro(AcqSynchro* p) { synchro_ = p; }
    void set_threshold(float t) { c_.threshold = t; }
    void set_doppler_max(uint32_t v)
    {
        c_.doppler_max = v;
        init();
    }
    void set_doppler_step(uint32_t v)
    {
        c_.doppler_step = v;
        init();
    }
    // = init() (galileo_pcps_8ms_acquisition_cc.cc:140-172): float(GALILEO_TWO_PI) * doppler / float(fs), all in float32
    void init()
    {
        const double GALILEO_TWO_PI = 6.283185307179600;
        bins_ = detail::inclusive_bins(c_.doppler_max, c_.doppler_step);
        std::vector<float> steps(bins_);
        for (uint32_t i = 0; i < bins_; i++)
            {
                const int32_t doppler = -static_cast<int32_t>(c_.doppler_max) + static_cast<int32_t>(c_.doppler_step * i);
                const float phase_step_rad = static_cast<float>(GALILEO_TWO_PI) * doppler / static_cast<float>(c_.fs_in);
                steps[i] = phase_step_rad;
            }
        for (auto* h : h_) ok_ = ok_ && gb200_acq_set_doppler_grid(h, bins_, steps.data()) == GB200_OK;
    }
    // = set_local_code (:117-138): code A as given; code B = the same with its second code period multiplied by (-1, 0)
    bool set_local_code(const std::complex<float>* code, size_t n_items)
    {
        std::vector<std::complex<float>> b(code, code + n_items);
        const size_t spc = static_cast<size_t>(c_.samples_per_code);
        for (size_t i = spc; i < 2 * spc && i < n_items; i++) b[i] = code[i] * std::complex<float>(-1.0f, 0.0f);
        return gb200_acq_set_local_code(h_[0], reinterpret_cast<const float*>(code), n_items) == GB200_OK &&
               gb200_acq_set_local_code(h_[1], reinterpret_cast<const float*>(b.data()), n_items) == GB200_OK;
    }
    void set_state(int state)
    {
        state_ = state;
        if (state == 1) well_count_ = 0;
    }
    void reset() { set_state(1); }
    uint32_t fft_size() const { return gb200_acq_fft_size(h_[0]); }
    uint32_t num_doppler_bins() const { return bins_; }
    float test_statistics() const { return test_statistics_; }
    float mag() const { return mag_; }
    float input_power() const { return input_power_; }
    uint32_t code_phase_index() const { return indext_; }
    uint32_t doppler_bin() const { return bin_; }

    // one pass through general_work state 1 (:225-375); both codes share the wiped-off forward spectra in ONE batch.
    // Per bin the block keeps the larger of (A, B), A on ties, and over the bins the first strictly larger value; on the
    // two per-code winners that is: A unless B is larger, or equal in an earlier bin.
    int run(uint64_t samp_count)
    {
        if (state_ != 1) return 0;
        gb200_acq_result r[2];
        if (gb200_acq_run_batch(h_, 2, stream_, samp_count, r) != GB200_OK) return -1;
        well_count_++;
        const float fnf = static_cast<float>(fft_size()) * static_cast<float>(fft_size());
        const float m[2] = {r[0].peak_mag / (fnf * fnf), r[1].peak_mag / (fnf * fnf)};
        const int w = (m[1] > m[0] || (m[1] == m[0] && r[1].doppler_bin < r[0].doppler_bin)) ? 1 : 0;
        mag_ = m[w];
        indext_ = r[w].code_phase_index;
        bin_ = r[w].doppler_bin;
        input_power_ = r[0].input_power;
        if (synchro_)
            {
                synchro_->Acq_delay_samples = static_cast<double>(indext_ % static_cast<uint32_t>(c_.samples_per_code));
                synchro_->Acq_doppler_hz = static_cast<double>(-static_cast<int32_t>(c_.doppler_max) + static_cast<int32_t>(c_.doppler_step * bin_));
                synchro_->Acq_samplestamp_samples = r[w].acq_samplestamp;
                synchro_->Acq_doppler_step = c_.doppler_step;
            }
        test_statistics_ = mag_ / input_power_;
        int state = 1;
        if (test_statistics_ > c_.threshold)
            state = 2;
        else if (well_count_ == c_.max_dwells)
            state = 3;
        if (state == 1) return 0;
        state_ = 0;
        if (synchro_) synchro_->Flag_valid_acquisition = (state == 2);
        return state == 2 ? 1 : 2;
    }

private:
    gb200_service* s_;
    AcqVariantConf c_;
    gb200_acq* h_[2] = {nullptr, nullptr};
    AcqSynchro* synchro_ = nullptr;
    int stream_;
    bool ok_ = false;
    int state_ = 0;
    uint32_t bins_ = 0, well_count_ = 0, indext_ = 0, bin_ = 0;
    float mag_ = 0.0f, input_power_ = 0.0f, test_statistics_ = 0.0f;
};
// ------------------------------------------------------------------------------------------------ Galileo E5a non-coherent I/Q + CAF
// galileo_e5a_noncoherentIQ_acquisition_caf_cc: up to four code spectra (I_A, Q_A and -- for more than one code period --
// the B replicas whose first period is inverted) correlated against the same wiped-off spectra; per bin the stronger of
// A / B is kept per component, the Q row is added onto the I row (non-coherent combination) and the row maximum taken.
// The device does all of that per bin (one batch + gb200_acq_iq_combine); the scalar bookkeeping per bin, the CAF filter
// over the per-bin peak values and the decision stay here.
namespace detail
{
// :685-768.  Integer types as in the block: `d` is int in the first and last segment and unsigned in the middle one, and
// several distances are converted to unsigned before (or without) abs() -- bins beyond the centre then weigh hugely
// negative.  That is the reference's arithmetic and is kept.
inline void caf_filter(const float* vi, const float* vq, float* cv, unsigned int nb, int window_hz, unsigned int doppler_step)
{
    const int H = window_hz / (2 * static_cast<int>(doppler_step));
    const float wf = 0.5 / static_cast<float>(H);
    auto smooth = [&](const float* v, int lo, int hi, auto weight, float divisor) {
        float acc = 0;
        for (int i = lo; i < hi; i++) acc += v[i] * weight(i);
        acc /= divisor;
        return acc;
    };
    for (int d = 0; d < H; d++)
        {
            const float div = 1 + H + d - wf * H * (H + 1) / 2 - wf * d * (d + 1) / 2;
            cv[d] = smooth(vi, 0, H + d + 1, [&](int i) { return 1 - wf * static_cast<unsigned int>(d - i); }, div);
            if (vq) cv[d] += smooth(vq, 0, H + d + 1, [&](int i) { return 1 - wf * static_cast<unsigned int>(std::abs(d - i)); }, div);
        }
    for (unsigned int d = H; d < nb - H; d++)
        {
            const float div = 1 + 2 * H - 2 * wf * H * (H + 1) / 2;
            auto w = [&](int i) { return 1 - wf * static_cast<unsigned int>(d - i); };
            cv[d] = smooth(vi, static_cast<int>(d - H), static_cast<int>(d + H + 1), w, div);
            if (vq) cv[d] += smooth(vq, static_cast<int>(d - H), static_cast<int>(d + H + 1), w, div);
        }
    for (int d = nb - H; d < static_cast<int>(nb); d++)
        {
            const float div = 1 + H + (nb - d - 1) - wf * H * (H + 1) / 2 - wf * (nb - d - 1) * (nb - d) / 2;
            auto w = [&](int i) { return 1 - wf * (std::abs(d - i)); };
            cv[d] = smooth(vi, d - H, static_cast<int>(nb), w, div);
            if (vq) cv[d] += smooth(vq, d - H, static_cast<int>(nb), w, div);
        }
}
}  // namespace detail

class B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf
{
public:
    B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf(gb200_service* s, const AcqVariantConf& c, int stream_id) : s_(s), c_(c), stream_(stream_id)
    {
        // :95-104: Zero_padding forces the block's d_sampled_ms to 1 (no B replicas); d_fft_size keeps sampled_ms
        eff_sampled_ms_ = c.zero_padding > 0 ? 1u : c.sampled_ms;
        const gb200_acq_cfg g = detail::variant_cfg(c, false);
        const bool want[4] = {true, c.both_signal_components, eff_sampled_ms_ > 1, c.both_signal_components && eff_sampled_ms_ > 1};
        ok_ = true;
        for (int k = 0; k < 4 && ok_; k++)
            if (want[k]) ok_ = gb200_acq_create(s_, &g, &h_[k]) == GB200_OK && gb200_acq_keep_grid(h_[k], 1) == GB200_OK;
        if (ok_) init();
    }
    ~B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf()
    {
        for (auto* h : h_)
            if (h) gb200_acq_destroy(h);
    }
    B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf(const B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf&) = delete;
    B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf& operator=(const B200_Galileo_E5a_NoncoherentIQ_Acquisition_Caf&) = delete;
    bool ok() const { return ok_; }
    void set_gnss_synchro(AcqSynchro* p) { synchro_ = p; }
    void set_threshold(float t) { c_.threshold = t; }
    void set_doppler_max(uint32_t v)
    {
        c_.doppler_max = v;
        init();
    }
    void set_doppler_step(uint32_t v)
    {
        c_.doppler_step = v;
        init();
    }
    // = init() (:275-331): GALILEO_TWO_PI (double) * doppler / float(fs), rounded once to float32
    void init()
    {
        const double GALILEO_TWO_PI = 6.283185307179600;
        bins_ = detail::inclusive_bins(c_.doppler_max, c_.doppler_step);
        std::vector<float> steps(bins_);
        for (uint32_t i = 0; i < bins_; i++)
            {
                const int doppler = -static_cast<int>(c_.doppler_max) + static_cast<int>(c_.doppler_step * i);
                const float phase_step_rad = GALILEO_TWO_PI * doppler / static_cast<float>(c_.fs_in);
                steps[i] = phase_step_rad;
            }
        for (auto* h : h_)
            if (h) ok_ = ok_ && gb200_acq_set_doppler_grid(h, bins_, steps.data()) == GB200_OK;
        caf_.assign(bins_, 0.0f);
        caf_i_.assign(bins_, 0.0f);
        caf_q_.assign(bins_, 0.0f);
        rec_.resize(bins_);
    }
    // = set_local_code(codeI, codeQ) (:238-272).  The block builds the B replicas by overwriting the first code period of
    // the FFT input buffer, which still holds the previous transform's input: the tail of I_B is codeQ's when both
    // components are used.
    bool set_local_code(const std::complex<float>* codeI, const std::complex<float>* codeQ, size_t n_items)
    {
        std::vector<std::complex<float>> inbuf(codeI, codeI + n_items);
        bool ok = gb200_acq_set_local_code(h_[0], reinterpret_cast<const float*>(inbuf.data()), n_items) == GB200_OK;
        if (h_[1])
            {
                inbuf.assign(codeQ, codeQ + n_items);
                ok = ok && gb200_acq_set_local_code(h_[1], reinterpret_cast<const float*>(inbuf.data()), n_items) == GB200_OK;
            }
        const size_t spc = std::min(static_cast<size_t>(c_.samples_per_code), n_items);
        for (int k = 2; k < 4; k++)
            {
                if (!h_[k]) continue;
                const std::complex<float>* src = k == 2 ? codeI : codeQ;
                for (size_t i = 0; i < spc; i++) inbuf[i] = src[i] * std::complex<float>(-1.0f, 0.0f);
                ok = ok && gb200_acq_set_local_code(h_[k], reinterpret_cast<const float*>(inbuf.data()), n_items) == GB200_OK;
            }
        return ok;
    }
    void set_state(int state)
    {
        state_ = state;
        if (state == 1)
            {
                // state 0 with d_active (:385-404)
                well_count_ = 0;
                test_statistics_ = 0.0f;
                mag_ = 0.0f;
                input_power_ = 0.0f;
            }
    }
    void reset() { set_state(1); }
    uint32_t fft_size() const { return gb200_acq_fft_size(h_[0]); }
    uint32_t num_doppler_bins() const { return bins_; }
    float test_statistics() const { return test_statistics_; }
    float mag() const { return mag_; }
    float input_power() const { return input_power_; }
    uint32_t code_phase_index() const { return indext_; }
    uint32_t doppler_bin() const { return bin_; }
    int doppler_hz() const { return doppler_hz_; }
    bool updated() const { return updated_; }
    uint32_t well_count() const { return well_count_; }
    const std::vector<float>& caf_vector() const { return caf_; }
    const std::vector<gb200_acq_iq_bin>& bin_records() const { return rec_; }

    // one dwell = state 2 of general_work (:440-792).  Returns 0 (next dwell), 1 (positive), 2 (negative), -1 (engine error).
    int run(uint64_t samp_count)
    {
        if (state_ != 1) return 0;
        gb200_acq* list[4];
        int n = 0;
        for (auto* h : h_)
            if (h) list[n++] = h;
        gb200_acq_result r[4];
        if (gb200_acq_run_batch(list, n, stream_, samp_count, r) != GB200_OK) return -1;
        const float fnf = static_cast<float>(fft_size()) * static_cast<float>(fft_size());
        if (gb200_acq_iq_combine(h_[0], h_[1], h_[2], h_[3], fnf * fnf, rec_.data()) != GB200_OK) return -1;
        well_count_++;
        input_power_ = r[0].input_power;
        mag_ = 0.0f;
        updated_ = false;
        for (uint32_t b = 0; b < bins_; b++)
            {
                const gb200_acq_iq_bin& q = rec_[b];
                if (c_.caf_window_hz > 0)
                    {
                        caf_i_[b] = q.peak[(q.picked & 1u) ? 2 : 0];
                        if (h_[1]) caf_q_[b] = q.peak[(q.picked & 2u) ? 3 : 1];
                    }
                const float magt = q.mag / (fnf * fnf);
                if (mag_ < magt)
                    {
                        mag_ = magt;
                        if (test_statistics_ < (mag_ / input_power_) || !c_.bit_transition_flag)
                            {
                                indext_ = q.indext;
                                bin_ = b;
                                doppler_hz_ = -static_cast<int>(c_.doppler_max) + static_cast<int>(c_.doppler_step * b);
                                updated_ = true;
                                test_statistics_ = mag_ / input_power_;
                                if (synchro_)
                                    {
                                        synchro_->Acq_delay_samples = static_cast<double>(indext_ % static_cast<uint32_t>(c_.samples_per_code));
                                        synchro_->Acq_doppler_hz = static_cast<double>(doppler_hz_);
                                        synchro_->Acq_samplestamp_samples = r[0].acq_samplestamp;
                                        synchro_->Acq_doppler_step = c_.doppler_step;
                                    }
                            }
                    }
            }
        if (c_.caf_window_hz > 0)
            {
                detail::caf_filter(caf_i_.data(), h_[1] ? caf_q_.data() : nullptr, caf_.data(), bins_, c_.caf_window_hz, c_.doppler_step);
                uint32_t t = 0;  // volk_gnsssdr_32f_index_max_32u: first maximum
                for (uint32_t i = 1; i < bins_; i++)
                    if (caf_[i] > caf_[t]) t = i;
                doppler_hz_ = -static_cast<int>(c_.doppler_max) + static_cast<int>(c_.doppler_step * t);
                if (synchro_) synchro_->Acq_doppler_hz = static_cast<double>(doppler_hz_);
            }
        if (well_count_ != c_.max_dwells) return 0;
        const bool positive = test_statistics_ > c_.threshold;
        state_ = 0;
        if (synchro_) synchro_->Flag_valid_acquisition = positive;
        return positive ? 1 : 2;
    }

private:
    gb200_service* s_;
    AcqVariantConf c_;
    gb200_acq* h_[4] = {nullptr, nullptr, nullptr, nullptr};  // I_A, Q_A, I_B, Q_B
    AcqSynchro* synchro_ = nullptr;
    int stream_;
    bool ok_ = false, updated_ = false;
    int state_ = 0, doppler_hz_ = 0;
    uint32_t eff_sampled_ms_ = 1, bins_ = 0, well_count_ = 0, indext_ = 0, bin_ = 0;
    float mag_ = 0.0f, input_power_ = 0.0f, test_statistics_ = 0.0f;
    std::vector<float> caf_, caf_i_, caf_q_;
    std::vector<gb200_acq_iq_bin> rec_;
};
// ------------------------------------------------------------------------------------------------ fine Doppler
// pcps_acquisition_fine_doppler_cc: the grid is the plain sum of |.|^2 over max_dwells dwells, the statistic is the ratio of its
// maximum to the second peak of the same Doppler row (one chip either side excluded), and a positive acquisition refines the
// Doppler with an 8x zero-padded FFT of 10 ms of code-wiped samples.  The grid, both peak searches, the wipe-off and the long
// FFT run on the device; the state sequence, the replica (incl. the block's off-by-one std::rotate) and the bin-to-Hz table
// arithmetic stay here.
class B200_Pcps_Acquisition_Fine_Doppler
{
public:
    B200_Pcps_Acquisition_Fine_Doppler(gb200_service* s, const AcqVariantConf& c, int stream_id) : s_(s), c_(c), stream_(stream_id)
    {
        AcqVariantConf one = c;
        one.sampled_ms = 1;  // d_fft_size = d_samples_per_ms (:64-66)
        const gb200_acq_cfg g = detail::variant_cfg(one, true);
        ok_ = gb200_acq_create(s_, &g, &h_) == GB200_OK && gb200_acq_set_scaled_accumulation(h_, 2) == GB200_OK;
        if (ok_) set_doppler_step(c_.doppler_step);
    }
    ~B200_Pcps_Acquisition_Fine_Doppler()
    {
        if (h_) gb200_acq_destroy(h_);
    }
    B200_Pcps_Acquisition_Fine_Doppler(const B200_Pcps_Acquisition_Fine_Doppler&) = delete;
    B200_Pcps_Acquisition_Fine_Doppler& operator=(const B200_Pcps_Acquisition_Fine_Doppler&) = delete;
    bool ok() const { return ok_; }
    void set_gnss_synchro(AcqSynchro* p) { synchro_ = p; }
    void set_threshold(float t) { c_.threshold = t; }
    // = set_doppler_step + update_carrier_wipeoff (:146-160, :246-261): floor(|2 doppler_max| / step) points from -doppler_max,
    // phase step float(GPS_TWO_PI) * doppler / float(fs) in float32
    void set_doppler_step(unsigned int doppler_step)
    {
        const double GPS_TWO_PI = 6.283185307179586;
        c_.doppler_step = doppler_step;
        num_doppler_points_ = floor(std::abs(2 * static_cast<int>(c_.doppler_max)) / c_.doppler_step);
        std::vector<float> steps(num_doppler_points_);
        for (int i = 0; i < num_doppler_points_; i++)
            {
                const int doppler_hz = c_.doppler_step * i - static_cast<int>(c_.doppler_max);
                steps[i] = static_cast<float>(GPS_TWO_PI) * doppler_hz / static_cast<float>(c_.fs_in);
            }
        ok_ = ok_ && num_doppler_points_ > 0 && gb200_acq_set_doppler_grid(h_, num_doppler_points_, steps.data()) == GB200_OK;
    }
    void set_doppler_max(unsigned int doppler_max)  // (:210-217 of the header: stored; the grid follows at the next set_doppler_step)
    {
        c_.doppler_max = doppler_max;
        if (c_.doppler_step > 0) set_doppler_step(c_.doppler_step);
    }
    void init() {}  // (:208-219): only clears the Gnss_Synchro fields, which the caller owns here
    // = set_local_code (:190-195); the copy is what estimate_Doppler regenerates with gps_l1_ca_code_gen_complex_sampled (:415)
    bool set_local_code(const std::complex<float>* code, size_t n_items)
    {
        code_.assign(code, code + std::min<size_t>(n_items, fft_size()));
        return gb200_acq_set_local_code(h_, reinterpret_cast<const float*>(code), n_items) == GB200_OK;
    }
    void set_state(int state)
    {
        state_ = state;
        if (state == 1) gb200_acq_reset(h_);  // reset_grid (:232-243)
    }
    void reset() { set_state(1); }
    uint32_t fft_size() const { return gb200_acq_fft_size(h_); }
    int num_doppler_points() const { return num_doppler_points_; }
    float test_statistics() const { return test_statistics_; }
    float first_peak() const { return first_peak_; }
    float second_peak() const { return second_peak_; }
    uint32_t index_time() const { return index_time_; }
    uint32_t index_doppler() const { return index_doppler_; }
    uint32_t fine_index() const { return fine_index_; }
    bool fine_accepted() const { return fine_accepted_; }

    // states 1..5 of general_work (:520-650) for an attempt whose first item is `samp_count`; the stream must hold
    // max(max_dwells, 10) * fft_size items from there.  Returns 1 (positive), 2 (negative), -1 (engine error).
    int run(uint64_t samp_count)
    {
        if (state_ != 1) return 0;
        const uint32_t N = fft_size();
        gb200_acq_result r{};
        for (uint32_t d = 0; d < c_.max_dwells; d++)  // S1: compute_and_accumulate_grid
            if (gb200_acq_run(h_, stream_, samp_count + static_cast<uint64_t>(d) * N, &r) != GB200_OK) return -1;
        // S2: compute_CAF (:266-326)
        first_peak_ = r.peak_mag;
        index_time_ = r.code_phase_index;
        index_doppler_ = r.doppler_bin;
        const double GPS_L1_CA_CHIP_PERIOD = 9.7752e-07;
        const uint32_t samplesPerChip = ceil(GPS_L1_CA_CHIP_PERIOD * static_cast<float>(c_.fs_in));
        if (gb200_acq_second_peak(h_, index_doppler_, index_time_, samplesPerChip, &second_peak_) != GB200_OK) return -1;
        test_statistics_ = first_peak_ / second_peak_;
        double acq_doppler_hz = static_cast<double>(static_cast<int>(index_doppler_ * c_.doppler_step) - static_cast<int>(c_.doppler_max));
        const double acq_delay_samples = static_cast<double>(index_time_);
        fine_accepted_ = false;
        fine_index_ = 0;
        const bool positive = test_statistics_ > c_.threshold;
        if (positive)
            {
                // S3: estimate_Doppler (:400-479)
                const int zero_padding_factor = 8, prn_replicas = 10;
                const int signal_samples = prn_replicas * static_cast<int>(N);
                const int fft_size_extended = signal_samples * zero_padding_factor;
                std::vector<std::complex<float>> code_replica(signal_samples);
                std::copy(code_.begin(), code_.end(), code_replica.begin());
                const int shift_index = static_cast<int>(acq_delay_samples);
                if (shift_index != 0) std::rotate(code_replica.begin(), code_replica.begin() + (N - shift_index), code_replica.begin() + N - 1);
                for (int n = 0; n < prn_replicas - 1; n++) std::copy(code_replica.begin(), code_replica.begin() + N, code_replica.begin() + (n + 1) * N);
                float peak = 0.0f;
                if (gb200_acq_wiped_spectrum_peak(s_, stream_, samp_count, reinterpret_cast<const float*>(code_replica.data()), signal_samples, fft_size_extended,
                        &fine_index_, &peak) != GB200_OK)
                    return -1;
                // fftFreqBins[tmp_index_freq] (:443-459)
                const int k = static_cast<int>(fine_index_);
                float bin_hz;
                if (k < fft_size_extended / 2)
                    bin_hz = ((static_cast<float>(c_.fs_in) / 2.0) * static_cast<float>(k)) / (static_cast<float>(fft_size_extended) / 2.0);
                else
                    bin_hz = ((-static_cast<float>(c_.fs_in) / 2.0) * static_cast<float>(fft_size_extended / 2 - (k - fft_size_extended / 2))) /
                             (static_cast<float>(fft_size_extended) / 2.0);
                if (std::abs(bin_hz - acq_doppler_hz) < 1000)
                    {
                        acq_doppler_hz = static_cast<double>(bin_hz);
                        fine_accepted_ = true;
                    }
            }
        if (synchro_)
            {
                synchro_->Acq_delay_samples = acq_delay_samples;
                synchro_->Acq_doppler_hz = acq_doppler_hz;
                synchro_->Acq_samplestamp_samples = r.acq_samplestamp;
                synchro_->Acq_doppler_step = c_.doppler_step;
                synchro_->Flag_valid_acquisition = positive;
            }
        state_ = 0;
        return positive ? 1 : 2;
    }

private:
    gb200_service* s_;
    AcqVariantConf c_;
    gb200_acq* h_ = nullptr;
    AcqSynchro* synchro_ = nullptr;
    int stream_;
    bool ok_ = false, fine_accepted_ = false;
    int state_ = 0, num_doppler_points_ = 0;
    uint32_t index_time_ = 0, index_doppler_ = 0, fine_index_ = 0;
    float first_peak_ = 0.0f, second_peak_ = 0.0f, test_statistics_ = 0.0f;
    std::vector<std::complex<float>> code_;
};
}  // namespace gb200
