// TEST INFRASTRUCTURE ONLY.  The product's host-side DLL/PLL epoch loop (gnss-sdr-1_b200/host/gb200_tracking.hpp)
// instantiated with a CPU correlator made of the oracle's restatement of the reference multicorrelator
// (oracle/gnss_oracle.c go_multicorrelator_real <- cpu_multicorrelator_real_codes.cc:101-152).  Closed-loop parity tests
// run this next to the same loop on the B200 correlator and compare trajectories; on CPU it checks that the loop locks.
#include <complex>
#include <cstdint>
#include <memory>
#include <vector>

#include "../../gnss-sdr-1_b200/host/gb200_host_flat.h"
#include "../../gnss-sdr-1_b200/host/gb200_c_aid_tracking.hpp"
#include "../../gnss-sdr-1_b200/host/gb200_glonass_tracking.hpp"
#include "../../gnss-sdr-1_b200/host/gb200_kf_tracking.hpp"
#include "../../gnss-sdr-1_b200/host/gb200_tracking.hpp"
#include "../../oracle/gnss_oracle.h"

namespace
{
class OracleCorrelator
{
public:
    OracleCorrelator(const float* samples_iq, uint64_t n_samples, uint64_t first_index, bool high_dyn) : x_(samples_iq), n_(n_samples), first_(first_index), high_dyn_(high_dyn) {}
    bool set_local_code_and_taps(int code_len, const float* code, const float* shifts, int n_taps)
    {
        code_.assign(code, code + code_len);
        shifts_.assign(shifts, shifts + n_taps);
        return true;
    }
    bool set_taps(const float* shifts, int n_taps)
    {
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
            code_step, code_rate, high_dyn_ ? 1 : 0, reinterpret_cast<float*>(out));
        if (!data_.empty())
            {
                // correlator_data_cpu: one tap at the prompt's offset, identical NCO arguments (dll_pll_veml_tracking.cc:899-909)
                const float s0 = shifts_[prompt_];
                go_multicorrelator_real(x, n, data_.data(), (int)data_.size(), &s0, 1, rem_carr, phase_step, phase_rate, rem_code, code_step, code_rate, high_dyn_ ? 1 : 0,
                    reinterpret_cast<float*>(out + shifts_.size()));
            }
        return true;
    }

    // the six-argument overload (cpu_multicorrelator_real_codes.cc:155-170): code resampler as configured, plain rotator
    bool correlate(uint64_t first, float rem_carr, float phase_step, float rem_code, float code_step, float code_rate, int n, std::complex<float>* out)
    {
        if (first < first_ || first + (uint64_t)n > first_ + n_) return false;
        go_multicorrelator_real(x_ + 2 * (first - first_), n, code_.data(), (int)code_.size(), shifts_.data(), (int)shifts_.size(), rem_carr, phase_step, 0.0f, rem_code,
            code_step, code_rate, high_dyn_ ? 2 : 0, reinterpret_cast<float*>(out));
        return true;
    }

private:
    const float* x_;
    uint64_t n_, first_;
    std::vector<float> code_, data_, shifts_;
    int prompt_ = 0;
    bool high_dyn_ = false;
};

struct Channel
{
    std::unique_ptr<OracleCorrelator> corr;
    std::unique_ptr<gb200::DllPllVemlTracking<OracleCorrelator>> loop;
};
}  // namespace

extern "C" {
void* hol_open(const float* samples_iq, uint64_t n_samples, uint64_t first_index, const gbh_trk_conf* c)
{
    const gb200::DllPllConf d = gbh_to_conf(c);
    auto* ch = new Channel();
    ch->corr.reset(new OracleCorrelator(samples_iq, n_samples, first_index, d.high_dyn));
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
    gbh_fill_epoch(e, ch->loop->state(), out);
    return ok ? 0 : -2;
}
void hol_set_channel(void* h, unsigned channel) { static_cast<Channel*>(h)->loop->set_channel(channel); }
void hol_telemetry(void* h, int tlm_event) { static_cast<Channel*>(h)->loop->telemetry_to_trk(tlm_event); }
void hol_close(void* h) { delete static_cast<Channel*>(h); }
}

// ---- the GLONASS L1 C/A loop on the oracle's complex-code correlator (go_multicorrelator_cplx <- cpu_multicorrelator.cc:99-126) ----
namespace
{
class OracleCplxCorrelator
{
public:
    using out_t = std::complex<float>;
    OracleCplxCorrelator(const float* samples_iq, uint64_t n_samples, uint64_t first_index) : x_(samples_iq), n_(n_samples), first_(first_index) {}
    bool set_local_code_and_taps(int code_len, const std::complex<float>* code, const float* shifts, int n_taps)
    {
        code_.assign(reinterpret_cast<const float*>(code), reinterpret_cast<const float*>(code) + 2 * code_len);
        shifts_.assign(shifts, shifts + n_taps);
        return true;
    }
    bool correlate(uint64_t first, float rem_carr, float phase_step, float rem_code, float code_step, int n, std::complex<float>* out)
    {
        if (first < first_ || first + (uint64_t)n > first_ + n_) return false;
        go_multicorrelator_cplx(x_ + 2 * (first - first_), n, code_.data(), (int)code_.size() / 2, shifts_.data(), (int)shifts_.size(), rem_carr, phase_step, rem_code,
            code_step, reinterpret_cast<float*>(out));
        return true;
    }

private:
    const float* x_;
    uint64_t n_, first_;
    std::vector<float> code_, shifts_;
};
struct GloChannel
{
    std::unique_ptr<OracleCplxCorrelator> corr;
    std::unique_ptr<gb200::GlonassL1CaDllPllTracking<OracleCplxCorrelator>> loop;
};
}  // namespace

extern "C" {
void* hol_glo_open(const float* samples_iq, uint64_t n_samples, uint64_t first_index, const double* conf)
{
    gb200::GlonassTrkConf c;
    gbh_glo_conf(conf, &c);
    auto* ch = new GloChannel();
    ch->corr.reset(new OracleCplxCorrelator(samples_iq, n_samples, first_index));
    ch->loop.reset(new gb200::GlonassL1CaDllPllTracking<OracleCplxCorrelator>(c, ch->corr.get()));
    return ch;
}
int hol_glo_start(void* h, const float* code_iq, int freq_channel, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now)
{
    auto* ch = static_cast<GloChannel*>(h);
    ch->loop->set_sample_counter(now);
    return ch->loop->start_tracking(reinterpret_cast<const std::complex<float>*>(code_iq), freq_channel, acq_delay_samples, acq_doppler_hz, acq_samplestamp) ? 0 : -1;
}
int hol_glo_step(void* h, double* epoch)
{
    gb200::GlonassEpoch e;
    const bool ok = static_cast<GloChannel*>(h)->loop->step(&e);
    gbh_glo_fill(e, epoch);
    return ok ? 0 : -2;
}
void hol_glo_close(void* h) { delete static_cast<GloChannel*>(h); }
}

// ---- the carrier-aided GPS L1 C/A loops on the oracle's correlators: _cc on go_multicorrelator_cplx, _sc on go_multicorrelator_16sc ----
namespace
{
class Oracle16scCorrelator
{
public:
    using out_t = std::complex<int16_t>;
    Oracle16scCorrelator(const int16_t* samples_iq, uint64_t n_samples, uint64_t first_index) : x_(samples_iq), n_(n_samples), first_(first_index) {}
    // volk_gnsssdr_32fc_convert_16ic.h (generic): clamp, rintf
    bool set_local_code_and_taps(int code_len, const std::complex<float>* code, const float* shifts, int n_taps)
    {
        const float* f = reinterpret_cast<const float*>(code);
        code_.resize(2 * (size_t)code_len);
        for (size_t i = 0; i < code_.size(); i++)
            {
                float aux = f[i];
                if (aux > 32767.0f) aux = 32767.0f;
                if (aux < -32768.0f) aux = -32768.0f;
                code_[i] = (int16_t)rintf(aux);
            }
        shifts_.assign(shifts, shifts + n_taps);
        return true;
    }
    bool correlate(uint64_t first, float rem_carr, float phase_step, float rem_code, float code_step, int n, out_t* out)
    {
        if (first < first_ || first + (uint64_t)n > first_ + n_) return false;
        go_multicorrelator_16sc(x_ + 2 * (first - first_), n, code_.data(), (int)code_.size() / 2, shifts_.data(), (int)shifts_.size(), rem_carr, phase_step, rem_code,
            code_step, reinterpret_cast<int16_t*>(out));
        return true;
    }

private:
    const int16_t* x_;
    uint64_t n_, first_;
    std::vector<int16_t> code_;
    std::vector<float> shifts_;
};
struct CAidChannel
{
    int variant = 0;
    std::unique_ptr<OracleCplxCorrelator> corr_cc;
    std::unique_ptr<Oracle16scCorrelator> corr_sc;
    std::unique_ptr<gb200::GpsL1CaDllPllCAidTracking<OracleCplxCorrelator>> loop_cc;
    std::unique_ptr<gb200::GpsL1CaDllPllCAidTracking<Oracle16scCorrelator>> loop_sc;
};
}  // namespace

extern "C" {
// samples: float IQ pairs (variant 0) or int16 IQ pairs (variant 1)
void* hol_caid_open(const void* samples_iq, uint64_t n_samples, uint64_t first_index, int variant, const double* conf)
{
    gb200::CAidTrkConf c;
    gbh_caid_conf(conf, &c);
    auto* ch = new CAidChannel();
    ch->variant = variant;
    if (variant == 1)
        {
            ch->corr_sc.reset(new Oracle16scCorrelator(static_cast<const int16_t*>(samples_iq), n_samples, first_index));
            ch->loop_sc.reset(new gb200::GpsL1CaDllPllCAidTracking<Oracle16scCorrelator>(c, ch->corr_sc.get()));
        }
    else
        {
            ch->corr_cc.reset(new OracleCplxCorrelator(static_cast<const float*>(samples_iq), n_samples, first_index));
            ch->loop_cc.reset(new gb200::GpsL1CaDllPllCAidTracking<OracleCplxCorrelator>(c, ch->corr_cc.get()));
        }
    return ch;
}
int hol_caid_start(void* h, const float* code_iq, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint64_t now)
{
    auto* ch = static_cast<CAidChannel*>(h);
    const auto* code = reinterpret_cast<const std::complex<float>*>(code_iq);
    if (ch->variant == 1)
        {
            ch->loop_sc->set_sample_counter(now);
            return ch->loop_sc->start_tracking(code, acq_delay_samples, acq_doppler_hz, acq_samplestamp) ? 0 : -1;
        }
    ch->loop_cc->set_sample_counter(now);
    return ch->loop_cc->start_tracking(code, acq_delay_samples, acq_doppler_hz, acq_samplestamp) ? 0 : -1;
}
void hol_caid_set_preamble(void* h, double preamble_timestamp_s)
{
    auto* ch = static_cast<CAidChannel*>(h);
    if (ch->variant == 1)
        ch->loop_sc->set_preamble_timestamp(preamble_timestamp_s);
    else
        ch->loop_cc->set_preamble_timestamp(preamble_timestamp_s);
}
int hol_caid_step(void* h, double* epoch)
{
    auto* ch = static_cast<CAidChannel*>(h);
    gb200::CAidEpoch e;
    const bool ok = ch->variant == 1 ? ch->loop_sc->step(&e) : ch->loop_cc->step(&e);
    gbh_caid_fill(e, epoch);
    return ok ? 0 : -2;
}
void hol_caid_close(void* h) { delete static_cast<CAidChannel*>(h); }
}

// ---- the Kalman GPS L1 C/A loop (gb200_kf_tracking.hpp) on the oracle's real-code correlator ----
namespace
{
struct KfChannel
{
    std::unique_ptr<OracleCorrelator> corr;
    std::unique_ptr<gb200::GpsL1CaKfTracking<OracleCorrelator>> loop;
};
}  // namespace
extern "C" {
void* hol_kf_open(const float* samples_iq, uint64_t n_samples, uint64_t first_index, const double* conf)
{
    gb200::KfTrkConf c;
    gbh_kf_conf(conf, &c);
    auto* ch = new KfChannel();
    ch->corr.reset(new OracleCorrelator(samples_iq, n_samples, first_index, true));  // the constructor default of the reference class
    ch->loop.reset(new gb200::GpsL1CaKfTracking<OracleCorrelator>(c, ch->corr.get()));
    return ch;
}
int hol_kf_start(void* h, const float* code, double acq_delay_samples, double acq_doppler_hz, uint64_t acq_samplestamp, uint32_t acq_doppler_step, uint64_t now)
{
    auto* ch = static_cast<KfChannel*>(h);
    ch->loop->set_sample_counter(now);
    return ch->loop->start_tracking(code, acq_delay_samples, acq_doppler_hz, acq_samplestamp, acq_doppler_step) ? 0 : -1;
}
int hol_kf_step(void* h, double* epoch)
{
    gb200::KfEpoch e;
    const bool ok = static_cast<KfChannel*>(h)->loop->step(&e);
    gbh_kf_fill(e, epoch);
    return ok ? 0 : -2;
}
void hol_kf_close(void* h) { delete static_cast<KfChannel*>(h); }
}
