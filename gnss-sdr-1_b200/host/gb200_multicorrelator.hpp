// B200_Multicorrelator_Real_Codes: the class a maintainer drops in for Cpu_Multicorrelator_Real_Codes
// (src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.h:46-74) -- same methods, same argument meaning,
// bool returns -- implemented on the C ABI (include/gnss_b200.h).  sig_in is not a pointer here: the samples live in the
// service's device ring and a correlation names them by absolute sample index.
#pragma once
#include <complex>
#include <vector>
#include <cstdint>

#include "../../include/gnss_b200.h"

namespace gb200
{
class B200_Multicorrelator_Real_Codes
{
public:
    B200_Multicorrelator_Real_Codes() = default;
    ~B200_Multicorrelator_Real_Codes() { free(); }
    B200_Multicorrelator_Real_Codes(const B200_Multicorrelator_Real_Codes&) = delete;
    B200_Multicorrelator_Real_Codes& operator=(const B200_Multicorrelator_Real_Codes&) = delete;

    void set_high_dynamics_resampler(bool use) { high_dyn_ = use; }  // cpu_multicorrelator_real_codes.cc:55-59
    // = init (:62-76); the service replaces the per-object aligned scratch buffers
    bool init(gb200_service* service, int stream_id, int item_type, int max_signal_length_samples, int n_correlators)
    {
        free();
        s_ = service;
        stream_ = stream_id;
        n_ = n_correlators;
        max_len_ = max_signal_length_samples;
        gb200_trk_cfg c{};
        c.n_taps = n_correlators;
        c.max_code_len = 2 * 10230;
        c.code_kind = GB200_CODE_REAL;
        c.high_dyn = high_dyn_ ? 1 : 0;
        c.item_type = item_type;
        return gb200_trk_create(s_, &c, &h_) == GB200_OK;
    }
    // = set_local_code_and_taps (:79-89); shifts may be changed later (narrow spacing switch)
    bool set_local_code_and_taps(int code_length_chips, const float* local_code_in, const float* shifts_chips, int n_taps)
    {
        if (!h_ || n_taps != n_) return false;
        return gb200_trk_set_local_code_and_taps(h_, local_code_in, code_length_chips, shifts_chips, n_taps) == GB200_OK;
    }
    // the reference keeps a POINTER to the caller's shift array and the tracking block edits it in place when it switches to the
    // narrow spacings (dll_pll_veml_tracking.cc:1753-1764); here the new values are handed over explicitly
    bool set_taps(const float* shifts_chips, int n_taps)
    {
        if (!h_ || n_taps != n_) return false;
        return gb200_trk_set_taps(h_, shifts_chips, n_taps) == GB200_OK;
    }
    // pilot tracking: what correlator_data_cpu does with identical NCO arguments (dll_pll_veml_tracking.cc:899-909)
    bool set_data_code(const float* code, int code_length_chips, int prompt_tap)
    {
        if (!h_) return false;
        has_data_ = gb200_trk_set_data_code(h_, code, code_length_chips, prompt_tap) == GB200_OK;
        return has_data_;
    }
    // = set_input_output_vectors (:92-98) + Carrier_wipeoff_multicorrelator_resampler (:129-152)
    bool correlate(uint64_t first_sample, float rem_carrier_phase_in_rad, float phase_step_rad, float phase_rate_step_rad, float rem_code_phase_chips,
        float code_phase_step_chips, float code_phase_rate_step_chips, int signal_length_samples, std::complex<float>* corr_out)
    {
        if (!h_ || signal_length_samples > max_len_) return false;
        return gb200_trk_correlate(h_, stream_, first_sample, rem_carrier_phase_in_rad, phase_step_rad, phase_rate_step_rad, rem_code_phase_chips,
                   code_phase_step_chips, code_phase_rate_step_chips, signal_length_samples, reinterpret_cast<float*>(corr_out)) == GB200_OK;
    }
    // = the six-argument overload (:155-170): code resampler as configured, plain rotator (no carrier rate term)
    bool correlate(uint64_t first_sample, float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips,
        float code_phase_rate_step_chips, int signal_length_samples, std::complex<float>* corr_out)
    {
        return correlate(first_sample, rem_carrier_phase_in_rad, phase_step_rad, 0.0f, rem_code_phase_chips, code_phase_step_chips, code_phase_rate_step_chips,
            signal_length_samples, corr_out);
    }
    bool free()  // (:173-190)
    {
        if (h_) gb200_trk_destroy(h_);
        h_ = nullptr;
        return true;
    }
    gb200_trk* handle() const { return h_; }
    int n_outputs() const { return n_ + (has_data_ ? 1 : 0); }

private:
    gb200_service* s_ = nullptr;
    gb200_trk* h_ = nullptr;
    int stream_ = 0, n_ = 0, max_len_ = 0;
    bool high_dyn_ = false, has_data_ = false;
};
}  // namespace gb200

namespace gb200
{
// B200_Multicorrelator: the drop-in for Cpu_Multicorrelator (complex local codes, src/algorithms/tracking/libs/
// cpu_multicorrelator.h:47-70) that the GLONASS L1 / L2 C/A trackers own -- same methods and argument meaning.
class B200_Multicorrelator
{
public:
    using out_t = std::complex<float>;
    B200_Multicorrelator() = default;
    ~B200_Multicorrelator() { free(); }
    B200_Multicorrelator(const B200_Multicorrelator&) = delete;
    B200_Multicorrelator& operator=(const B200_Multicorrelator&) = delete;
    // = init (cpu_multicorrelator.cc:51-66)
    bool init(gb200_service* service, int stream_id, int item_type, int max_signal_length_samples, int n_correlators)
    {
        free();
        s_ = service;
        stream_ = stream_id;
        n_ = n_correlators;
        max_len_ = max_signal_length_samples;
        gb200_trk_cfg c{};
        c.n_taps = n_correlators;
        c.max_code_len = 2 * 10230;
        c.code_kind = GB200_CODE_CPLX;
        c.item_type = item_type;
        return gb200_trk_create(s_, &c, &h_) == GB200_OK;
    }
    // = set_local_code_and_taps (:69-79): local_code_in holds code_length_chips complex values
    bool set_local_code_and_taps(int code_length_chips, const std::complex<float>* local_code_in, const float* shifts_chips, int n_taps)
    {
        if (!h_ || n_taps != n_) return false;
        return gb200_trk_set_local_code_and_taps(h_, local_code_in, code_length_chips, shifts_chips, n_taps) == GB200_OK;
    }
    // = set_input_output_vectors (:82-88) + Carrier_wipeoff_multicorrelator_resampler (:99-126)
    bool correlate(uint64_t first_sample, float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips,
        int signal_length_samples, std::complex<float>* corr_out)
    {
        if (!h_ || signal_length_samples > max_len_) return false;
        return gb200_trk_correlate(h_, stream_, first_sample, rem_carrier_phase_in_rad, phase_step_rad, 0.0f, rem_code_phase_chips, code_phase_step_chips, 0.0f,
                   signal_length_samples, reinterpret_cast<float*>(corr_out)) == GB200_OK;
    }
    bool free()  // (:129-141)
    {
        if (h_) gb200_trk_destroy(h_);
        h_ = nullptr;
        return true;
    }

private:
    gb200_service* s_ = nullptr;
    gb200_trk* h_ = nullptr;
    int stream_ = 0, n_ = 0, max_len_ = 0;
};

// = Cpu_Multicorrelator_16sc (src/algorithms/tracking/libs/cpu_multicorrelator_16sc.h:47-72): lv_16sc_t samples (a
// GB200_ITEM_16IC stream), lv_16sc_t local code, lv_16sc_t outputs -- bit-exact against the reference's generic kernels.
class B200_Multicorrelator_16sc
{
public:
    using out_t = std::complex<int16_t>;
    B200_Multicorrelator_16sc() = default;
    ~B200_Multicorrelator_16sc() { free(); }
    B200_Multicorrelator_16sc(const B200_Multicorrelator_16sc&) = delete;
    B200_Multicorrelator_16sc& operator=(const B200_Multicorrelator_16sc&) = delete;
    // = init (cpu_multicorrelator_16sc.cc:49-66)
    bool init(gb200_service* service, int stream_id, int max_signal_length_samples, int n_correlators)
    {
        free();
        s_ = service;
        stream_ = stream_id;
        n_ = n_correlators;
        max_len_ = max_signal_length_samples;
        gb200_trk_cfg c{};
        c.n_taps = n_correlators;
        c.max_code_len = 2 * 10230;
        c.code_kind = GB200_CODE_16SC;
        c.item_type = GB200_ITEM_16IC;
        return gb200_trk_create(s_, &c, &h_) == GB200_OK;
    }
    // The trackers hand over a gr_complex code and convert it first (volk_gnsssdr_32fc_convert_16ic, e.g.
    // gps_l1_ca_dll_pll_c_aid_tracking_sc.cc:252-255): clamp to the int16 range, round to nearest even (rintf)
    bool set_local_code_and_taps(int code_length_chips, const std::complex<float>* local_code_in, const float* shifts_chips, int n_taps)
    {
        if (!h_ || n_taps != n_) return false;
        std::vector<int16_t> c16(2 * static_cast<size_t>(code_length_chips));
        const float* f = reinterpret_cast<const float*>(local_code_in);
        for (size_t i = 0; i < c16.size(); i++)
            {
                float aux = f[i];
                if (aux > 32767.0f)
                    aux = 32767.0f;
                else if (aux < -32768.0f)
                    aux = -32768.0f;
                c16[i] = static_cast<int16_t>(rintf(aux));
            }
        return gb200_trk_set_local_code_and_taps(h_, c16.data(), code_length_chips, shifts_chips, n_taps) == GB200_OK;
    }
    // = set_input_output_vectors + Carrier_wipeoff_multicorrelator_resampler (:78-105)
    bool correlate(uint64_t first_sample, float rem_carrier_phase_in_rad, float phase_step_rad, float rem_code_phase_chips, float code_phase_step_chips,
        int signal_length_samples, out_t* corr_out)
    {
        if (!h_ || signal_length_samples > max_len_) return false;
        return gb200_trk_correlate_16sc(h_, stream_, first_sample, rem_carrier_phase_in_rad, phase_step_rad, rem_code_phase_chips, code_phase_step_chips,
                   signal_length_samples, reinterpret_cast<int16_t*>(corr_out)) == GB200_OK;
    }
    bool free()
    {
        if (h_) gb200_trk_destroy(h_);
        h_ = nullptr;
        return true;
    }

private:
    gb200_service* s_ = nullptr;
    gb200_trk* h_ = nullptr;
    int stream_ = 0, n_ = 0, max_len_ = 0;
};
}  // namespace gb200
