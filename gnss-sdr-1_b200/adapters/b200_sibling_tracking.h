// b200_sibling_tracking<Policy>: ONE GNU Radio block template for the older per-signal trackers of the reference, each of which
// is its own gr::block there:
//   glonass_l1_ca_dll_pll_tracking_cc, glonass_l2_ca_dll_pll_tracking_cc (src/algorithms/tracking/gnuradio_blocks/glonass_l{1,2}_ca_dll_pll_tracking_cc.cc)
//   gps_l1_ca_dll_pll_c_aid_tracking_cc / _sc (.../gps_l1_ca_dll_pll_c_aid_tracking_cc.cc, _sc.cc)
//   gps_l1_ca_kf_tracking_cc                 (.../gps_l1_ca_kf_tracking_cc.cc)
// The loops are the host mirrors of gnss-sdr-1_b200/host/ (each pinned to its block) on the B200 correlators; a
// Policy names the loop, its correlator, how the local code is made and how start_tracking is called.  What the blocks share is
// here: one Gnss_Synchro out per general_work (also in standby and for the pull-in alignment call), consume = the loop's block
// length, event 3 on loss of lock, forecast = 2 * vector_length, the "preamble_timestamp_s" port of the carrier-aided trackers.
#ifndef GNSS_SDR_B200_SIBLING_TRACKING_H
#define GNSS_SDR_B200_SIBLING_TRACKING_H

#include "GLONASS_L1_L2_CA.h"
#include "GPS_L1_CA.h"
#include "glonass_l1_signal_processing.h"
#include "glonass_l2_signal_processing.h"
#include "gnss_synchro.h"
#include "gps_sdr_signal_processing.h"
#include "host/gb200_c_aid_tracking.hpp"
#include "host/gb200_glonass_tracking.hpp"
#include "host/gb200_kf_tracking.hpp"
#include "host/gb200_multicorrelator.hpp"
#include <gnss_b200.h>
#include <gnuradio/block.h>
#include <gnuradio/io_signature.h>
#include <pmt/pmt.h>
#include <boost/shared_ptr.hpp>
#include <complex>
#include <cstring>
#include <memory>
#include <vector>

// ---- policies --------------------------------------------------------------------------------------------------------------
struct B200GlonassL1Policy
{
    using Conf = gb200::GlonassTrkConf;
    using Corr = gb200::B200_Multicorrelator;
    using Loop = gb200::GlonassL1CaDllPllTracking<Corr>;
    using Epoch = gb200::GlonassEpoch;
    static const char* name() { return "b200_glonass_l1_ca_dll_pll_tracking"; }
    static char system() { return 'R'; }
    static bool init_corr(Corr& c, gb200_service* s, int stream, int item_type, int vlen) { return c.init(s, stream, item_type, 2 * vlen, 3); }
    static bool start(Loop& l, const Gnss_Synchro& syn)
    {
        // glonass_l1_ca_dll_pll_tracking_cc.cc:257: glonass_l1_ca_code_gen_complex(d_ca_code, 0); channel number of the slot :223
        std::vector<std::complex<float>> code(static_cast<size_t>(GLONASS_L1_CA_CODE_LENGTH_CHIPS));
        glonass_l1_ca_code_gen_complex(code.data(), 0);
        const int k = GLONASS_PRN.count(syn.PRN) ? GLONASS_PRN.at(syn.PRN) : 0;
        return l.start_tracking(code.data(), k, syn.Acq_delay_samples, syn.Acq_doppler_hz, syn.Acq_samplestamp_samples);
    }
    static void preamble(Loop&, double) {}
    static int consumed(const Epoch& e) { return e.prn_length_samples; }
    static int length_ms(const Epoch&) { return 1; }
};

struct B200GlonassL2Policy : B200GlonassL1Policy
{
    // glonass_l2_ca_dll_pll_tracking_cc.cc: the same loop with the L2 band constants (GlonassTrkConf::band_freq_hz / band_dfreq_hz) and generator
    static const char* name() { return "b200_glonass_l2_ca_dll_pll_tracking"; }
    static bool start(Loop& l, const Gnss_Synchro& syn)
    {
        std::vector<std::complex<float>> code(static_cast<size_t>(GLONASS_L2_CA_CODE_LENGTH_CHIPS));
        glonass_l2_ca_code_gen_complex(code.data(), 0);
        const int k = GLONASS_PRN.count(syn.PRN) ? GLONASS_PRN.at(syn.PRN) : 0;
        return l.start_tracking(code.data(), k, syn.Acq_delay_samples, syn.Acq_doppler_hz, syn.Acq_samplestamp_samples);
    }
};

template <class CorrT>
struct B200CAidPolicyBase
{
    using Conf = gb200::CAidTrkConf;
    using Corr = CorrT;
    using Loop = gb200::GpsL1CaDllPllCAidTracking<Corr>;
    using Epoch = gb200::CAidEpoch;
    static char system() { return 'G'; }
    static bool start(Loop& l, const Gnss_Synchro& syn)
    {
        // gps_l1_ca_dll_pll_c_aid_tracking_cc.cc:246: gps_l1_ca_code_gen_complex(d_ca_code, PRN, 0) (the _sc block converts it to int16, _sc.cc:252-255)
        std::vector<std::complex<float>> code(static_cast<size_t>(GPS_L1_CA_CODE_LENGTH_CHIPS));
        gps_l1_ca_code_gen_complex(code.data(), syn.PRN, 0);
        return l.start_tracking(code.data(), syn.Acq_delay_samples, syn.Acq_doppler_hz, syn.Acq_samplestamp_samples);
    }
    static void preamble(Loop& l, double t) { l.set_preamble_timestamp(t); }
    static int consumed(const Epoch& e) { return e.correlation_length_samples; }
    static int length_ms(const Epoch& e) { return e.correlation_length_ms; }
};
struct B200CAidCcPolicy : B200CAidPolicyBase<gb200::B200_Multicorrelator>
{
    static const char* name() { return "b200_gps_l1_ca_dll_pll_c_aid_tracking_cc"; }
    static bool init_corr(Corr& c, gb200_service* s, int stream, int item_type, int vlen) { return item_type == GB200_ITEM_32FC && c.init(s, stream, item_type, 2 * vlen, 3); }
};
struct B200CAidScPolicy : B200CAidPolicyBase<gb200::B200_Multicorrelator_16sc>
{
    static const char* name() { return "b200_gps_l1_ca_dll_pll_c_aid_tracking_sc"; }
    static bool init_corr(Corr& c, gb200_service* s, int stream, int item_type, int vlen) { return item_type == GB200_ITEM_16IC && c.init(s, stream, 2 * vlen, 3); }
};

struct B200KfPolicy
{
    using Conf = gb200::KfTrkConf;
    using Corr = gb200::B200_Multicorrelator_Real_Codes;
    using Loop = gb200::GpsL1CaKfTracking<Corr>;
    using Epoch = gb200::KfEpoch;
    static const char* name() { return "b200_gps_l1_ca_kf_tracking"; }
    static char system() { return 'G'; }
    static bool init_corr(Corr& c, gb200_service* s, int stream, int item_type, int vlen)
    {
        c.set_high_dynamics_resampler(true);  // the reference correlator's constructor default, which gps_l1_ca_kf_tracking_cc never changes
        return c.init(s, stream, item_type, 2 * vlen, 3);
    }
    static bool start(Loop& l, const Gnss_Synchro& syn)
    {
        std::vector<float> code(static_cast<size_t>(GPS_L1_CA_CODE_LENGTH_CHIPS));
        gps_l1_ca_code_gen_float(code.data(), syn.PRN, 0);  // gps_l1_ca_kf_tracking_cc.cc:324
        return l.start_tracking(code.data(), syn.Acq_delay_samples, syn.Acq_doppler_hz, syn.Acq_samplestamp_samples, syn.Acq_doppler_step);
    }
    static void preamble(Loop&, double) {}
    static int consumed(const Epoch& e) { return e.prn_length_samples; }
    static int length_ms(const Epoch&) { return 1; }
};

// ---- the block -------------------------------------------------------------------------------------------------------------
template <class Policy>
class b200_sibling_tracking : public gr::block
{
public:
    using sptr = boost::shared_ptr<b200_sibling_tracking<Policy>>;
    static sptr make(gb200_service* service, int stream_id, const typename Policy::Conf& conf, int item_type)
    {
        return sptr(new b200_sibling_tracking<Policy>(service, stream_id, conf, item_type));
    }
    ~b200_sibling_tracking() = default;

    void set_channel(uint32_t channel) { d_channel = channel; }
    void set_gnss_synchro(Gnss_Synchro* p) { d_acquisition_gnss_synchro = p; }
    bool ok() const { return ok_; }
    void start_tracking()
    {
        gr::thread::scoped_lock l(d_setlock);
        loop_->set_sample_counter(d_sample_counter);
        started_ = ok_ && d_acquisition_gnss_synchro && Policy::start(*loop_, *d_acquisition_gnss_synchro);
    }
    void stop_tracking() {}  // the reference blocks have no stop: a channel simply stops reading them

    void forecast(int noutput_items, gr_vector_int& ninput_items_required)
    {
        if (noutput_items != 0) ninput_items_required[0] = static_cast<int>(conf_.vector_length) * 2;
    }

    int general_work(int noutput_items __attribute__((unused)), gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
    {
        gr::thread::scoped_lock l(d_setlock);
        const char* in = reinterpret_cast<const char*>(input_items[0]);
        auto** out = reinterpret_cast<Gnss_Synchro**>(&output_items[0]);
        if (ninput_items[0] > 0) gb200_stream_push_shared(s_, stream_, in, static_cast<size_t>(ninput_items[0]), item_type_, d_sample_counter);
        typename Policy::Epoch e;
        const bool tracking = started_ && loop_->tracking_enabled();
        const bool stepped = loop_->step(&e);
        d_sample_counter = loop_->sample_counter();
        consume_each(Policy::consumed(e));
        Gnss_Synchro current_synchro_data = Gnss_Synchro();
        if (!stepped)
            {
                // the block's samples are no longer in the ring (this block lagged more than the ring holds): treated as loss of lock
                this->message_port_pub(pmt::mp("events"), pmt::from_long(3));
                started_ = false;
            }
        else if (tracking)
            {
                current_synchro_data = *d_acquisition_gnss_synchro;
                current_synchro_data.Prompt_I = e.prompt_i;
                current_synchro_data.Prompt_Q = e.prompt_q;
                current_synchro_data.Tracking_sample_counter = e.valid_symbol ? e.tracking_sample_counter : current_synchro_data.Tracking_sample_counter;
                current_synchro_data.Code_phase_samples = e.code_phase_samples;
                current_synchro_data.Carrier_phase_rads = e.carrier_phase_rads;
                current_synchro_data.Carrier_Doppler_hz = e.carrier_doppler_hz;
                current_synchro_data.CN0_dB_hz = e.cn0_db_hz;
                current_synchro_data.Flag_valid_symbol_output = e.valid_symbol;
                current_synchro_data.correlation_length_ms = Policy::length_ms(e) > 0 ? Policy::length_ms(e) : 1;
                if (e.loss_of_lock) this->message_port_pub(pmt::mp("events"), pmt::from_long(3));
            }
        else
            {
                current_synchro_data.System = Policy::system();
                current_synchro_data.Tracking_sample_counter = e.tracking_sample_counter;
                current_synchro_data.correlation_length_ms = 1;
            }
        current_synchro_data.fs = static_cast<int64_t>(conf_.fs_in);
        *out[0] = current_synchro_data;
        return 1;  // "output tracking result ALWAYS even in the case of d_enable_tracking==false"
    }

private:
    b200_sibling_tracking(gb200_service* service, int stream_id, const typename Policy::Conf& conf, int item_type)
        : gr::block(Policy::name(), gr::io_signature::make(1, 1, item_type == GB200_ITEM_32FC ? 8 : (item_type == GB200_ITEM_16IC ? 4 : 2)),
              gr::io_signature::make(1, 1, sizeof(Gnss_Synchro))),
          s_(service), stream_(stream_id), item_type_(item_type), conf_(conf)
    {
        this->message_port_register_out(pmt::mp("events"));
        this->message_port_register_in(pmt::mp("preamble_timestamp_s"));
        this->set_msg_handler(pmt::mp("preamble_timestamp_s"), boost::bind(&b200_sibling_tracking<Policy>::msg_handler_preamble_index, this, _1));
        this->message_port_register_in(pmt::mp("telemetry_to_trk"));
        this->set_relative_rate(1.0 / static_cast<double>(conf_.vector_length));
        ok_ = Policy::init_corr(corr_, s_, stream_, item_type_, static_cast<int>(conf_.vector_length));
        loop_.reset(new typename Policy::Loop(conf_, &corr_));
    }
    void msg_handler_preamble_index(pmt::pmt_t msg)  // gps_l1_ca_dll_pll_c_aid_tracking_cc.cc:90-107
    {
        gr::thread::scoped_lock l(d_setlock);
        Policy::preamble(*loop_, pmt::to_double(msg));
    }

    gb200_service* s_;
    int stream_, item_type_;
    typename Policy::Conf conf_;
    typename Policy::Corr corr_;
    std::unique_ptr<typename Policy::Loop> loop_;
    Gnss_Synchro* d_acquisition_gnss_synchro = nullptr;
    uint32_t d_channel = 0;
    uint64_t d_sample_counter = 0;
    bool ok_ = false, started_ = false;
};

#endif
