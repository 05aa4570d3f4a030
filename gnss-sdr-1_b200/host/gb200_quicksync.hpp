// B200_Pcps_QuickSync_Acquisition: host mirror of pcps_quicksync_acquisition_cc
// (src/algorithms/acquisition/gnuradio_blocks/pcps_quicksync_acquisition_cc.cc; adapter
// src/algorithms/acquisition/adapters/gps_l1_ca_pcps_quicksync_acquisition.cc) on the C ABI.
//
// The block searches a FOLDED signal: of every input vector (sampled_ms * samples_per_ms items) it takes samples_per_code *
// folding_factor items, wipes the Doppler off, adds the folding_factor^2 segments of fft_size = samples_per_code / folding_factor
// items onto each other and runs the PCPS correlation on those fft_size points (:380-425).  The maximum gives the delay modulo
// fft_size; the folding_factor candidates are then told apart by correlating one code period of the wiped, UNFOLDED signal with
// the unfolded code in the time domain (:436-470).
//   device (gb200_acq_create_folded): wipe-off + folding + FFT + x conj(code FFT) + IFFT + |.|^2 + arg-max for every bin, input power;
//   host   (here)                   : normalisation and statistic, well counter, threshold rule, and the candidate choice -- F
//                                     dot products of samples_per_code items on the winning bin only, on items read back from the
//                                     device ring (gb200_stream_read; 32fc streams, the only item type the block has).
// Two deliberate simplifications against the block's per-bin bookkeeping, both exact except on float ties:
//   * the block runs the candidate choice every time a bin improves d_mag and keeps the last one; only the last improvement -- the
//     dwell's first global maximum, which is what the device returns -- is observable, so it is run once;
//   * with bit_transition_flag the block also asks d_test_statistics < d_mag / d_input_power before it records an improvement; both
//     restart at 0 every dwell (:343-346) and the power is constant within one, so the test only fails when two successive maxima
//     round to the same quotient.
// The block adds the time-domain sums onto an uninitialised stack array (complex_acumulator[100], :439); here they start at 0.
#pragma once
#include <cmath>
#include <complex>
#include <cstdint>
#include <vector>

#include "../../include/gnss_b200.h"
#include "gb200_acquisition.hpp"

namespace gb200
{
struct QuickSyncConf
{
    uint32_t folding_factor = 2;
    uint32_t sampled_ms = 4;  // "coherent_integration_time_ms", a multiple of folding_factor (adapter :65-90)
    uint32_t max_dwells = 1;
    uint32_t doppler_max = 5000;
    uint32_t doppler_step = 250;
    int64_t fs_in = 4000000;
    int32_t samples_per_ms = 4000;
    int32_t samples_per_code = 4000;
    bool bit_transition_flag = false;  // the adapter then sets max_dwells = 2 (:95-102)
    float threshold = 0.0f;
};

class B200_Pcps_QuickSync_Acquisition
{
public:
    B200_Pcps_QuickSync_Acquisition(gb200_service* s, const QuickSyncConf& c, int stream_id) : s_(s), c_(c), stream_(stream_id)
    {
        fft_size_ = static_cast<uint32_t>(c_.samples_per_code) / c_.folding_factor;  // :95
        used_ = static_cast<uint32_t>(c_.samples_per_code) * c_.folding_factor;
        gb200_acq_cfg g{};
        g.sampled_ms = 1;  // one dwell reads the `used_` items at the head of the vector
        g.ms_per_code = 1;
        g.samples_per_chip = 1;
        g.max_dwells = 1;
        g.doppler_max = c_.doppler_max;
        g.doppler_step = c_.doppler_step ? c_.doppler_step : 250;
        g.fs_in = c_.fs_in;
        g.samples_per_ms = static_cast<float>(used_);
        g.samples_per_code = static_cast<float>(c_.samples_per_code);
        g.use_CFAR_algorithm_flag = 1;
        g.item_type = GB200_ITEM_32FC;
        g.threshold = c_.threshold;
        ok_ = fft_size_ * c_.folding_factor == static_cast<uint32_t>(c_.samples_per_code) && c_.folding_factor >= 2 && c_.folding_factor <= 100 &&
              gb200_acq_create_folded(s_, &g, c_.folding_factor * c_.folding_factor, &h_) == GB200_OK;
        if (ok_) init();
    }
    ~B200_Pcps_QuickSync_Acquisition()
    {
        if (h_) gb200_acq_destroy(h_);
    }
    B200_Pcps_QuickSync_Acquisition(const B200_Pcps_QuickSync_Acquisition&) = delete;
    B200_Pcps_QuickSync_Acquisition& operator=(const B200_Pcps_QuickSync_Acquisition&) = delete;
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
    // = init() (:201-243): inclusive bin count; float phase_step_rad = GPS_TWO_PI * doppler / float(fs) (double product, rounded once)
    void init()
    {
        const double GPS_TWO_PI = 6.283185307179586;
        if (c_.doppler_step == 0) c_.doppler_step = 250;
        bins_ = 0;
        for (auto d = -static_cast<int32_t>(c_.doppler_max); d <= static_cast<int32_t>(c_.doppler_max); d += static_cast<int32_t>(c_.doppler_step)) bins_++;
        steps_.resize(bins_);
        for (uint32_t i = 0; i < bins_; i++)
            {
                const int32_t doppler = -static_cast<int32_t>(c_.doppler_max) + static_cast<int32_t>(c_.doppler_step * i);
                const float phase_step_rad = GPS_TWO_PI * doppler / static_cast<float>(c_.fs_in);
                steps_[i] = phase_step_rad;
            }
        ok_ = ok_ && gb200_acq_set_doppler_grid(h_, bins_, steps_.data()) == GB200_OK;
    }
    // = set_local_code (:177-200): the unfolded code is kept for the time-domain step; folding_factor segments are added onto a
    // zeroed buffer (in that order, float32) and handed to the device, which transforms and conjugates them
    bool set_local_code(const std::complex<float>* code, size_t n_items)
    {
        if (n_items < static_cast<size_t>(c_.samples_per_code)) return false;
        code_.assign(code, code + c_.samples_per_code);
        std::vector<std::complex<float>> folded(fft_size_, std::complex<float>(0.0f, 0.0f));
        for (uint32_t i = 0; i < c_.folding_factor; i++)
            for (uint32_t k = 0; k < fft_size_; k++) folded[k] = code[static_cast<size_t>(i) * fft_size_ + k] + folded[k];
        return gb200_acq_set_local_code(h_, reinterpret_cast<const float*>(folded.data()), fft_size_) == GB200_OK;
    }
    // = set_state(1) (:246-270) / state 0 with d_active (:296-310)
    void set_state(int state)
    {
        state_ = state;
        if (state == 1)
            {
                if (synchro_)
                    {
                        synchro_->Acq_delay_samples = 0.0;
                        synchro_->Acq_doppler_hz = 0.0;
                        synchro_->Acq_samplestamp_samples = 0ULL;
                        synchro_->Acq_doppler_step = 0U;
                    }
                well_count_ = 0;
                mag_ = 0.0f;
                input_power_ = 0.0f;
                test_statistics_ = 0.0f;
                gb200_acq_reset(h_);
            }
    }
    void reset() { set_state(1); }
    uint32_t fft_size() const { return fft_size_; }
    uint32_t vector_items() const { return c_.sampled_ms * static_cast<uint32_t>(c_.samples_per_ms); }  // what one dwell consumes (consume_each(1))
    uint32_t num_doppler_bins() const { return bins_; }
    uint32_t well_count() const { return well_count_; }
    float mag() const { return mag_; }
    float test_statistics() const { return test_statistics_; }
    float input_power() const { return input_power_; }
    uint32_t folded_index() const { return indext_; }
    const std::vector<float>& corr_output() const { return corr_output_f_; }  // d_corr_output_f of the last dwell

    // one pass through general_work state 1 (:318-532) on the input vector that starts at samp_count.
    // Returns the event: 1 positive, 2 negative, 0 keep dwelling (call again with the next vector), -1 error.
    int run(uint64_t samp_count)
    {
        if (state_ != 1) return 0;
        gb200_acq_result r;
        if (gb200_acq_run(h_, stream_, samp_count, &r) != GB200_OK) return -1;
        const float fft_normalization_factor = static_cast<float>(fft_size_) * static_cast<float>(fft_size_);
        // :343-349
        mag_ = 0.0f;
        test_statistics_ = 0.0f;
        const uint64_t sample_counter = samp_count + vector_items();  // d_sample_counter after :349
        well_count_++;
        input_power_ = r.input_power;
        const float magt = r.peak_mag / (fft_normalization_factor * fft_normalization_factor);  // :421
        if (mag_ < magt)
            {
                mag_ = magt;
                indext_ = r.code_phase_index;
                const int32_t doppler = -static_cast<int32_t>(c_.doppler_max) + static_cast<int32_t>(c_.doppler_step * r.doppler_bin);
                uint32_t best = 0;
                if (!choose_candidate(samp_count, r.doppler_bin, indext_ % static_cast<uint32_t>(c_.samples_per_code), &best)) return -1;
                if (synchro_)
                    {
                        synchro_->Acq_delay_samples = static_cast<double>(possible_delay_[best]);
                        synchro_->Acq_doppler_hz = static_cast<double>(doppler);
                        synchro_->Acq_samplestamp_samples = sample_counter;
                        synchro_->Acq_doppler_step = c_.doppler_step;
                    }
                test_statistics_ = mag_ / input_power_;  // :479
            }
        // :503-527
        int state = 1;
        if (!c_.bit_transition_flag)
            {
                if (test_statistics_ > c_.threshold)
                    state = 2;
                else if (well_count_ == c_.max_dwells)
                    state = 3;
            }
        else if (well_count_ == c_.max_dwells)
            state = test_statistics_ > c_.threshold ? 2 : 3;
        if (state == 1) return 0;
        state_ = 0;  // states 2 / 3 (:537-600): d_active = false, d_state = 0, event published
        if (synchro_) synchro_->Flag_valid_acquisition = (state == 2);
        return state == 2 ? 1 : 2;
    }

private:
    // :436-470 for Doppler bin `bin`: candidate i starts at folded + i * fft_size; correlate samples_per_code wiped items with the
    // unfolded code, keep the first maximum of |.|^2.  The wipe-off table entry n is (cosf, sinf) of a float32 phase that starts at
    // 0 and is advanced by -phase_step_rad n times (volk_gnsssdr_s32f_sincos_32fc), so it is accumulated from 0 here as well.
    bool choose_candidate(uint64_t samp_count, uint32_t bin, uint32_t folded, uint32_t* best)
    {
        const uint32_t F = c_.folding_factor, spc = static_cast<uint32_t>(c_.samples_per_code);
        const uint32_t need = folded + (F - 1) * fft_size_ + spc;  // <= used_
        if (code_.size() != spc || need > used_) return false;
        in_.resize(need);
        if (gb200_stream_read(s_, stream_, samp_count, need, in_.data()) != GB200_OK) return false;
        const float phase_inc = -steps_[bin];
        float phase = 0.0f;
        for (uint32_t n = 0; n < need; n++)
            {
                in_[n] = in_[n] * std::complex<float>(cosf(phase), sinf(phase));  // volk_32fc_x2_multiply_32fc(in_temp, in, wipeoff)
                phase += phase_inc;
            }
        possible_delay_.resize(F);
        corr_output_f_.resize(F);
        for (uint32_t i = 0; i < F; i++)
            {
                possible_delay_[i] = folded + i * fft_size_;
                std::complex<float> acc(0.0f, 0.0f);
                const std::complex<float>* x = in_.data() + possible_delay_[i];
                for (uint32_t j = 0; j < spc; j++) acc += x[j] * code_[j];
                corr_output_f_[i] = acc.real() * acc.real() + acc.imag() * acc.imag();
            }
        *best = 0;
        for (uint32_t i = 1; i < F; i++)
            if (corr_output_f_[i] > corr_output_f_[*best]) *best = i;  // volk_gnsssdr_32f_index_max_32u: first maximum
        return true;
    }

    gb200_service* s_;
    QuickSyncConf c_;
    gb200_acq* h_ = nullptr;
    AcqSynchro* synchro_ = nullptr;
    int stream_;
    bool ok_ = false;
    int state_ = 0;
    uint32_t fft_size_ = 0, used_ = 0, bins_ = 0, well_count_ = 0, indext_ = 0;
    float mag_ = 0.0f, input_power_ = 0.0f, test_statistics_ = 0.0f;
    std::vector<float> steps_, corr_output_f_;
    std::vector<std::complex<float>> code_, in_;
    std::vector<uint32_t> possible_delay_;
};
}  // namespace gb200
