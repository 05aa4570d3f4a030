// TEST INFRASTRUCTURE ONLY.  The product's host-side DLL/PLL epoch loop (gnss-sdr-1_b200/host/gb200_tracking.hpp)
// instantiated with a CPU correlator made of the oracle's restatement of the reference multicorrelator
// (oracle/gnss_oracle.c go_multicorrelator_real <- cpu_multicorrelator_real_codes.cc:101-152).  Closed-loop parity tests
// run this next to the same loop on the B200 correlator and compare trajectories; on CPU it checks that the loop locks.
#include <complex>
#include <cstdint>
#include <memory>
#include <vector>

#include "../../gnss-sdr-1_b200/host/gb200_tracking.hpp"
#include "../../oracle/gnss_oracle.h"

namespace
{
class OracleCorrelator
{
public:
    OracleCorrelator(const float* samples_iq, uint64_t n_samples, uint64_t first_index) : x_(samples_iq), n_(n_samples), first_(first_index) {}
    bool set_local_code_and_taps(int code_len, const float* code, const float* shifts, int n_taps)
    {
        code_.assign(code, code + code_len);
        shifts_.assign(shifts, shifts + n_taps);
        return true;
    }
    bool set_data_code(const float* code, int code_len, int prompt_tap)
    {
        data_.assign(code, code + code_len);
        prompt_ = prompt_tap;
        return true;
    }
    bool correlate(uint64_t first, float rem_carr, float phase_step, float phase_rate, float rem_code, float code_step, float code_rate, int n,
        std::complex<float>* out)
    {
        if (first < first_ || first + (uint64_t)n > first_ + n_) return false;
        const float* x = x_ + 2 * (first - first_);
        go_multicorrelator_real(x, n, code_.data(), (int)code_.size(), shifts_.data(), (int)shifts_.size(), rem_carr, phase_step, phase_rate, rem_code,
            code_step, code_rate, 0, reinterpret_cast<float*>(out));
        if (!data_.empty())
            {
                // correlator_data_cpu: one tap at the prompt's offset, identical NCO arguments (dll_pll_veml_tracking.cc:899-909)
                const float s0 = shifts_[prompt_];
                go_multicorrelator_real(x, n, data_.data(), (int)data_.size(), &s0, 1, rem_carr, phase_step, phase_rate, rem_code, code_step, code_rate, 0,
                    reinterpret_cast<float*>(out + shifts_.size()));
            }
        return true;
    }

private:
    const float* x_;
    uint64_t n_, first_;
    std::vector<float> code_, data_, shifts_;
    int prompt_ = 0;
};

struct Channel
{
    std::unique_ptr<OracleCorrelator> corr;
    std::unique_ptr<gb200::DllPllVemlTracking<OracleCorrelator>> loop;
};
}  // namespace

// the flat structs of gnss-sdr-1_b200/host/host_abi.cc (same layout: the Python side shares one ctypes definition)
struct gbh_trk_conf
{
    double fs_in;
    int vector_length;
    float pll_bw_hz, dll_bw_hz, fll_bw_hz;
    int pll_filter_order, dll_filter_order;
    int enable_fll_pull_in, enable_fll_steady_state;
    float pull_in_time_s, early_late_space_chips, very_early_late_space_chips;
    int cn0_samples;
    float cn0_min;
    int max_lock_fail;
    float carrier_lock_th;
    int track_pilot, veml;
    double signal_carrier_freq, code_chip_rate;
    int code_length_chips, code_samples_per_chip;
    double code_period;
};
struct gbh_epoch
{
    int valid, loss_of_lock, state, current_prn_length_samples;
    uint64_t tracking_sample_counter;
    double prompt_i, prompt_q, code_phase_samples, carrier_phase_rads, carrier_doppler_hz, cn0_db_hz, code_error_chips, carr_phase_error_hz, code_freq_chips;
    float taps_iq[12];
};

extern "C" {
void* hol_open(const float* samples_iq, uint64_t n_samples, uint64_t first_index, const gbh_trk_conf* c)
{
    gb200::DllPllConf d;
    d.fs_in = c->fs_in;
    d.vector_length = c->vector_length;
    d.pll_bw_hz = c->pll_bw_hz;
    d.dll_bw_hz = c->dll_bw_hz;
    d.fll_bw_hz = c->fll_bw_hz;
    d.pll_filter_order = c->pll_filter_order;
    d.dll_filter_order = c->dll_filter_order;
    d.enable_fll_pull_in = c->enable_fll_pull_in != 0;
    d.enable_fll_steady_state = c->enable_fll_steady_state != 0;
    d.pull_in_time_s = c->pull_in_time_s;
    d.early_late_space_chips = c->early_late_space_chips;
    d.very_early_late_space_chips = c->very_early_late_space_chips;
    d.cn0_samples = c->cn0_samples;
    d.cn0_min = c->cn0_min;
    d.max_lock_fail = c->max_lock_fail;
    d.carrier_lock_th = c->carrier_lock_th;
    d.track_pilot = c->track_pilot != 0;
    d.veml = c->veml != 0;
    d.signal_carrier_freq = c->signal_carrier_freq;
    d.code_chip_rate = c->code_chip_rate;
    d.code_length_chips = c->code_length_chips;
    d.code_samples_per_chip = c->code_samples_per_chip;
    d.code_period = c->code_period;
    auto* ch = new Channel();
    ch->corr.reset(new OracleCorrelator(samples_iq, n_samples, first_index));
    ch->loop.reset(new gb200::DllPllVemlTracking<OracleCorrelator>(d, ch->corr.get()));
    return ch;
}
int hol_start(void* h, const float* code, const float* data_code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now)
{
    return static_cast<Channel*>(h)->loop->start_tracking(code, data_code, acq_delay_samples, acq_doppler_hz, acq_samplestamp, now) ? 0 : -1;
}
int hol_step(void* h, gbh_epoch* out)
{
    auto* ch = static_cast<Channel*>(h);
    gb200::TrackingEpoch e;
    const bool ok = ch->loop->step(&e);
    out->valid = e.flag_valid_symbol_output;
    out->loss_of_lock = e.loss_of_lock;
    out->state = ch->loop->state();
    out->current_prn_length_samples = e.current_prn_length_samples;
    out->tracking_sample_counter = e.tracking_sample_counter;
    out->prompt_i = e.prompt_i;
    out->prompt_q = e.prompt_q;
    out->code_phase_samples = e.code_phase_samples;
    out->carrier_phase_rads = e.carrier_phase_rads;
    out->carrier_doppler_hz = e.carrier_doppler_hz;
    out->cn0_db_hz = e.cn0_db_hz;
    out->code_error_chips = e.code_error_chips;
    out->carr_phase_error_hz = e.carr_phase_error_hz;
    out->code_freq_chips = e.code_freq_chips;
    for (int t = 0; t < 6; t++)
        {
            out->taps_iq[2 * t] = e.taps[t].real();
            out->taps_iq[2 * t + 1] = e.taps[t].imag();
        }
    return ok ? 0 : -2;
}
void hol_close(void* h) { delete static_cast<Channel*>(h); }
}
