// Flat POD images of DllPllConf / TrackingEpoch for ctypes, shared by libgnss_b200_host.so (host_abi.cc) and the test-side build
// of the same epoch loop on a CPU correlator (tests/native/).  tests/host_lib.py holds the matching ctypes definitions.
#pragma once
#include <cstdint>
#include <cstring>

#include "gb200_tracking.hpp"

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
    // narrow tracking / extended integration / high dynamics / per-signal constants
    float pll_bw_narrow_hz, dll_bw_narrow_hz, early_late_space_narrow_chips, very_early_late_space_narrow_chips;
    int extend_correlation_symbols, high_dyn, smoother_length;
    int secondary, symbols_per_bit, interchange_iq, correlation_length_ms;
    uint32_t prn;
    int n_preamble_symbols;
    int32_t preamble_symbols[200];
    char secondary_code[128];   // NUL-terminated '0'/'1' string
    char dump_filename[256];    // NUL-terminated, empty = no dump
};

struct gbh_epoch
{
    int valid, loss_of_lock, state, current_prn_length_samples;
    uint64_t tracking_sample_counter;
    double prompt_i, prompt_q, code_phase_samples, carrier_phase_rads, carrier_doppler_hz, cn0_db_hz, code_error_chips, carr_phase_error_hz, code_freq_chips;
    float taps_iq[12];
    int correlation_length_ms, reserved_;
};

inline gb200::DllPllConf gbh_to_conf(const gbh_trk_conf* c)
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
    d.pll_bw_narrow_hz = c->pll_bw_narrow_hz;
    d.dll_bw_narrow_hz = c->dll_bw_narrow_hz;
    d.early_late_space_narrow_chips = c->early_late_space_narrow_chips;
    d.very_early_late_space_narrow_chips = c->very_early_late_space_narrow_chips;
    d.extend_correlation_symbols = c->extend_correlation_symbols;
    d.high_dyn = c->high_dyn != 0;
    d.smoother_length = c->smoother_length;
    d.secondary = c->secondary != 0;
    d.symbols_per_bit = c->symbols_per_bit;
    d.interchange_iq = c->interchange_iq != 0;
    d.correlation_length_ms = c->correlation_length_ms;
    d.prn = c->prn;
    const int np = c->n_preamble_symbols < 0 ? 0 : (c->n_preamble_symbols > 200 ? 200 : c->n_preamble_symbols);
    d.preamble_symbols.assign(c->preamble_symbols, c->preamble_symbols + np);
    d.secondary_code.assign(c->secondary_code, strnlen(c->secondary_code, sizeof c->secondary_code));
    d.dump_filename.assign(c->dump_filename, strnlen(c->dump_filename, sizeof c->dump_filename));
    return d;
}

inline void gbh_fill_epoch(const gb200::TrackingEpoch& e, int state, gbh_epoch* out)
{
    out->valid = e.flag_valid_symbol_output;
    out->loss_of_lock = e.loss_of_lock;
    out->state = state;
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
    out->correlation_length_ms = e.correlation_length_ms;
    out->reserved_ = 0;
}
