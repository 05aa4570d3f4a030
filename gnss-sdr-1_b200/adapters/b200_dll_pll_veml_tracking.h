// b200_dll_pll_veml_tracking: the GNU Radio block behind the *_DLL_PLL_*_Tracking_B200 adapters -- the public surface of
// dll_pll_veml_tracking (src/algorithms/tracking/gnuradio_blocks/dll_pll_veml_tracking.h:53-75: set_channel, set_gnss_synchro,
// start_tracking, stop_tracking, one Gnss_Synchro out per code period, "events" 3 = loss of lock, "telemetry_to_trk" fault in).
// The epoch loop is gb200::DllPllVemlTracking (host/gb200_tracking.hpp, bit-identical to the reference block on the reference's
// correlator: tests/test_ref_blocks_cpu.py) on B200_Multicorrelator_Real_Codes, i.e. every do_correlation_step is one
// gb200_trk_correlate that the shared service coalesces with the other channels' epochs into one launch.
#ifndef GNSS_SDR_B200_DLL_PLL_VEML_TRACKING_H
#define GNSS_SDR_B200_DLL_PLL_VEML_TRACKING_H

#include "Beidou_B1I.h"
#include "Beidou_B3I.h"
#include "GPS_L2C.h"
#include "GPS_L1_CA.h"
#include "GPS_L5.h"
#include "Galileo_E1.h"
#include "Galileo_E5a.h"
#include "beidou_b1i_signal_processing.h"
#include "beidou_b3i_signal_processing.h"
#include "gps_l2c_signal.h"
#include "dll_pll_conf.h"
#include "galileo_e1_signal_processing.h"
#include "galileo_e5_signal_processing.h"
#include "gnss_synchro.h"
#include "gps_l5_signal.h"
#include "gps_sdr_signal_processing.h"
#include "host/gb200_multicorrelator.hpp"
#include "host/gb200_tracking.hpp"
#include <gnss_b200.h>
#include <gnuradio/block.h>
#include <gnuradio/io_signature.h>
#include <pmt/pmt.h>
#include <complex>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

class b200_dll_pll_veml_tracking;
using b200_dll_pll_veml_tracking_sptr = boost::shared_ptr<b200_dll_pll_veml_tracking>;

class b200_dll_pll_veml_tracking : public gr::block
{
public:
    static b200_dll_pll_veml_tracking_sptr make(gb200_service* service, int stream_id, const Dll_Pll_Conf& conf_, int item_type)
    {
        return b200_dll_pll_veml_tracking_sptr(new b200_dll_pll_veml_tracking(service, stream_id, conf_, item_type));
    }
    ~b200_dll_pll_veml_tracking() = default;

    void set_channel(uint32_t channel)
    {
        gr::thread::scoped_lock l(d_setlock);
        d_channel = channel;
        if (loop_) loop_->set_channel(channel);
    }
    void set_gnss_synchro(Gnss_Synchro* p_gnss_synchro)
    {
        gr::thread::scoped_lock l(d_setlock);
        d_acquisition_gnss_synchro = p_gnss_synchro;
    }
    void stop_tracking()
    {
        gr::thread::scoped_lock l(d_setlock);
        if (loop_) loop_->stop_tracking();
    }

    // = start_tracking (:549-747): the local code tables come from the reference's own generators, exactly as there
    void start_tracking()
    {
        gr::thread::scoped_lock l(d_setlock);
        const Gnss_Synchro& syn = *d_acquisition_gnss_synchro;
        gb200::DllPllConf c = base_;
        c.prn = syn.PRN;
        const std::string sig(trk_parameters.signal);
        std::vector<float> code(2 * c.code_length_chips), data(2 * c.code_length_chips);
        bool have_data = false;
        if (trk_parameters.system == 'G' && sig == "1C")
            gps_l1_ca_code_gen_float(code.data(), syn.PRN, 0);
        else if (trk_parameters.system == 'G' && sig == "2S")
            gps_l2c_m_code_gen_float(code.data(), syn.PRN);  // (:586-589)
        else if (trk_parameters.system == 'G' && sig == "L5")
            {
                if (trk_parameters.track_pilot)
                    {
                        gps_l5q_code_gen_float(code.data(), syn.PRN);
                        gps_l5i_code_gen_float(data.data(), syn.PRN);
                        have_data = true;
                    }
                else
                    gps_l5i_code_gen_float(code.data(), syn.PRN);
            }
        else if (trk_parameters.system == 'E' && sig == "1B")
            {
                char pilot_signal[3] = "1C";
                char data_signal[3] = "1B";
                if (trk_parameters.track_pilot)
                    {
                        galileo_e1_code_gen_sinboc11_float(code.data(), pilot_signal, syn.PRN);
                        galileo_e1_code_gen_sinboc11_float(data.data(), data_signal, syn.PRN);
                        have_data = true;
                    }
                else
                    galileo_e1_code_gen_sinboc11_float(code.data(), data_signal, syn.PRN);
            }
        else if (trk_parameters.system == 'E' && sig == "5X")
            {
                std::vector<std::complex<float>> aux(c.code_length_chips);
                char s5x[3] = "5X";
                galileo_e5_a_code_gen_complex_primary(aux.data(), syn.PRN, s5x);
                if (trk_parameters.track_pilot)
                    {
                        c.secondary_code = GALILEO_E5A_Q_SECONDARY_CODE[syn.PRN - 1];
                        for (int i = 0; i < c.code_length_chips; i++)
                            {
                                code[i] = aux[i].imag();
                                data[i] = aux[i].real();
                            }
                        have_data = true;
                    }
                else
                    for (int i = 0; i < c.code_length_chips; i++) code[i] = aux[i].real();
            }
        else if (trk_parameters.system == 'C' && sig == "B1")
            {
                beidou_b1i_code_gen_float(code.data(), syn.PRN, 0);
                if (syn.PRN > 0 && syn.PRN < 6)
                    {
                        // GEO satellites (:630-667): no secondary code, D2 preamble of 11 bits x 2 symbols
                        c.symbols_per_bit = 2;
                        c.secondary = false;
                        c.secondary_code.clear();
                        c.preamble_symbols.clear();
                        const int bits[11] = {1, 1, 1, 0, 0, 0, 1, 0, 0, 1, 0};
                        for (int b : bits)
                            for (int j = 0; j < 2; j++) c.preamble_symbols.push_back(b ? 1 : -1);
                    }
            }
        else if (trk_parameters.system == 'C' && sig == "B3")
            {
                beidou_b3i_code_gen_float(code.data(), syn.PRN, 0);
                if (syn.PRN > 0 && syn.PRN < 6)
                    {
                        // GEO satellites (:668-705): no secondary code, D2 preamble of 11 bits x 2 symbols
                        c.symbols_per_bit = 2;
                        c.secondary = false;
                        c.secondary_code.clear();
                        c.preamble_symbols.clear();
                        const int bits[11] = {1, 1, 1, 0, 0, 0, 1, 0, 0, 1, 0};
                        for (int b : bits)
                            for (int j = 0; j < 2; j++) c.preamble_symbols.push_back(b ? 1 : -1);
                    }
            }
        loop_.reset(new gb200::DllPllVemlTracking<gb200::B200_Multicorrelator_Real_Codes>(c, &corr_));
        loop_->set_channel(d_channel);
        started_ = loop_->start_tracking(code.data(), have_data ? data.data() : nullptr, syn.Acq_delay_samples, syn.Acq_doppler_hz, syn.Acq_samplestamp_samples,
            d_sample_counter);
    }

    void forecast(int noutput_items, gr_vector_int& ninput_items_required)
    {
        if (noutput_items != 0) ninput_items_required[0] = static_cast<int32_t>(trk_parameters.vector_length) * 2;  // (:507-514)
    }

    int general_work(int noutput_items __attribute__((unused)), gr_vector_int& ninput_items, gr_vector_const_void_star& input_items, gr_vector_void_star& output_items)
    {
        gr::thread::scoped_lock l(d_setlock);
        const char* in = reinterpret_cast<const char*>(input_items[0]);
        auto** out = reinterpret_cast<Gnss_Synchro**>(&output_items[0]);
        // whatever the scheduler shows this block is offered to the shared ring first (the first block to see an item copies it)
        if (ninput_items[0] > 0) gb200_stream_push_shared(s_, stream_, in, static_cast<size_t>(ninput_items[0]), item_type_, d_sample_counter);
        if (!loop_ || !started_ || loop_->state() == 0)
            {
                // standby (:1562-1567): consume at full throttle
                d_sample_counter += static_cast<uint64_t>(ninput_items[0]);
                consume_each(ninput_items[0]);
                return 0;
            }
        gb200::TrackingEpoch e;
        const uint64_t before = loop_->sample_counter();
        const bool ok = loop_->step(&e);
        const uint64_t consumed = loop_->sample_counter() - before;
        d_sample_counter = loop_->sample_counter();
        consume_each(static_cast<int>(consumed));
        if (!ok)
            {
                // the epoch's samples are no longer in the ring (this block lagged more than it holds): treated as loss of lock
                loop_->stop_tracking();
                this->message_port_pub(pmt::mp("events"), pmt::from_long(3));
                return 0;
            }
        if (e.loss_of_lock) this->message_port_pub(pmt::mp("events"), pmt::from_long(3));  // (:873)
        if (!e.flag_valid_symbol_output) return 0;
        // states 3 / 4 start from the acquisition's Gnss_Synchro (System, Signal, PRN, Acq_*), state 2 from a blank one (:1775, :1829)
        Gnss_Synchro current_synchro_data = narrow_or_extended_ ? *d_acquisition_gnss_synchro : Gnss_Synchro();
        narrow_or_extended_ = loop_->state() >= 3;
        current_synchro_data.Prompt_I = e.prompt_i;
        current_synchro_data.Prompt_Q = e.prompt_q;
        current_synchro_data.Code_phase_samples = e.code_phase_samples;
        current_synchro_data.Carrier_phase_rads = e.carrier_phase_rads;
        current_synchro_data.Carrier_Doppler_hz = e.carrier_doppler_hz;
        current_synchro_data.CN0_dB_hz = e.cn0_db_hz;
        current_synchro_data.Flag_valid_symbol_output = true;
        current_synchro_data.correlation_length_ms = e.correlation_length_ms;
        current_synchro_data.fs = static_cast<int64_t>(trk_parameters.fs_in);
        current_synchro_data.Tracking_sample_counter = e.tracking_sample_counter;
        *out[0] = current_synchro_data;
        return 1;
    }
    int state() const { return loop_ ? loop_->state() : 0; }

private:
    b200_dll_pll_veml_tracking(gb200_service* service, int stream_id, const Dll_Pll_Conf& conf_, int item_type)
        : gr::block("b200_dll_pll_veml_tracking", gr::io_signature::make(1, 1, item_type == GB200_ITEM_32FC ? 8 : (item_type == GB200_ITEM_16IC ? 4 : 2)),
              gr::io_signature::make(1, 1, sizeof(Gnss_Synchro))),
          s_(service), stream_(stream_id), item_type_(item_type), trk_parameters(conf_)
    {
        this->message_port_register_out(pmt::mp("events"));
        this->set_relative_rate(1.0 / static_cast<double>(trk_parameters.vector_length));
        this->message_port_register_in(pmt::mp("telemetry_to_trk"));
        this->set_msg_handler(pmt::mp("telemetry_to_trk"), boost::bind(&b200_dll_pll_veml_tracking::msg_handler_telemetry_to_trk, this, _1));
        derive_signal_constants();
        corr_.set_high_dynamics_resampler(trk_parameters.high_dyn);
        corr_.init(s_, stream_, item_type_, 2 * static_cast<int>(trk_parameters.vector_length), base_.veml ? 5 : 3);
    }

    void msg_handler_telemetry_to_trk(const pmt::pmt_t& msg)  // (:516-546)
    {
        try
            {
                if (pmt::any_ref(msg).type() == typeid(int))
                    {
                        const int tlm_event = boost::any_cast<int>(pmt::any_ref(msg));
                        gr::thread::scoped_lock lock(d_setlock);
                        if (loop_) loop_->telemetry_to_trk(tlm_event);
                    }
            }
        catch (boost::bad_any_cast& e)
            {
            }
    }

    // what the reference's constructor derives from (system, signal) (:113-336)
    void derive_signal_constants()
    {
        gb200::DllPllConf& c = base_;
        const Dll_Pll_Conf& p = trk_parameters;
        c.fs_in = p.fs_in;
        c.vector_length = static_cast<int>(p.vector_length);
        c.pll_bw_hz = p.pll_bw_hz;
        c.dll_bw_hz = p.dll_bw_hz;
        c.fll_bw_hz = p.fll_bw_hz;
        c.pll_filter_order = p.pll_filter_order;
        c.dll_filter_order = p.dll_filter_order;
        c.enable_fll_pull_in = p.enable_fll_pull_in;
        c.enable_fll_steady_state = p.enable_fll_steady_state;
        c.pull_in_time_s = static_cast<float>(p.pull_in_time_s);
        c.early_late_space_chips = p.early_late_space_chips;
        c.very_early_late_space_chips = p.very_early_late_space_chips;
        c.early_late_space_narrow_chips = p.early_late_space_narrow_chips;
        c.very_early_late_space_narrow_chips = p.very_early_late_space_narrow_chips;
        c.pll_bw_narrow_hz = p.pll_bw_narrow_hz;
        c.dll_bw_narrow_hz = p.dll_bw_narrow_hz;
        c.extend_correlation_symbols = p.extend_correlation_symbols;
        c.high_dyn = p.high_dyn;
        c.smoother_length = static_cast<int>(p.smoother_length);
        c.cn0_samples = p.cn0_samples;
        c.cn0_min = static_cast<float>(p.cn0_min);
        c.max_lock_fail = p.max_lock_fail;
        c.carrier_lock_th = static_cast<float>(p.carrier_lock_th);
        c.track_pilot = p.track_pilot;
        if (p.dump) c.dump_filename = p.dump_filename;
        const std::string sig(p.signal);
        c.veml = false;
        c.interchange_iq = false;
        c.secondary = false;
        c.symbols_per_bit = 1;
        c.code_samples_per_chip = 1;
        c.correlation_length_ms = 1;
        if (p.system == 'G' && sig == "1C")
            {
                c.signal_carrier_freq = GPS_L1_FREQ_HZ;
                c.code_period = GPS_L1_CA_CODE_PERIOD;
                c.code_chip_rate = GPS_L1_CA_CODE_RATE_HZ;
                c.symbols_per_bit = GPS_CA_TELEMETRY_SYMBOLS_PER_BIT;
                c.code_length_chips = static_cast<int>(GPS_L1_CA_CODE_LENGTH_CHIPS);
                c.track_pilot = false;
                const uint16_t preambles_bits[GPS_CA_PREAMBLE_LENGTH_BITS] = GPS_PREAMBLE;
                for (uint16_t b : preambles_bits)
                    for (uint32_t j = 0; j < GPS_CA_TELEMETRY_SYMBOLS_PER_BIT; j++) c.preamble_symbols.push_back(b == 1 ? 1 : -1);
            }
        else if (p.system == 'G' && sig == "2S")
            {
                // (:155-168): 20 ms code, no pilot, no secondary code
                c.signal_carrier_freq = GPS_L2_FREQ_HZ;
                c.code_period = GPS_L2_M_PERIOD;
                c.code_chip_rate = GPS_L2_M_CODE_RATE_HZ;
                c.code_length_chips = static_cast<int>(GPS_L2_M_CODE_LENGTH_CHIPS);
                c.symbols_per_bit = GPS_L2_SAMPLES_PER_SYMBOL;
                c.correlation_length_ms = 20;
                c.track_pilot = false;
            }
        else if (p.system == 'G' && sig == "L5")
            {
                c.signal_carrier_freq = GPS_L5_FREQ_HZ;
                c.code_period = GPS_L5I_PERIOD;
                c.code_chip_rate = GPS_L5I_CODE_RATE_HZ;
                c.symbols_per_bit = GPS_L5_SAMPLES_PER_SYMBOL;
                c.code_length_chips = static_cast<int>(GPS_L5I_CODE_LENGTH_CHIPS);
                c.secondary = true;
                c.secondary_code = p.track_pilot ? GPS_L5Q_NH_CODE_STR : GPS_L5I_NH_CODE_STR;
                c.interchange_iq = p.track_pilot;
            }
        else if (p.system == 'E' && sig == "1B")
            {
                c.signal_carrier_freq = GALILEO_E1_FREQ_HZ;
                c.code_period = GALILEO_E1_CODE_PERIOD;
                c.code_chip_rate = GALILEO_E1_CODE_CHIP_RATE_HZ;
                c.code_length_chips = static_cast<int>(GALILEO_E1_B_CODE_LENGTH_CHIPS);
                c.symbols_per_bit = 1;
                c.correlation_length_ms = 4;
                c.code_samples_per_chip = 2;
                c.veml = true;
                if (p.track_pilot)
                    {
                        c.secondary = true;
                        c.secondary_code = GALILEO_E1_C_SECONDARY_CODE;
                    }
            }
        else if (p.system == 'E' && sig == "5X")
            {
                c.signal_carrier_freq = GALILEO_E5A_FREQ_HZ;
                c.code_period = GALILEO_E5A_CODE_PERIOD;
                c.code_chip_rate = GALILEO_E5A_CODE_CHIP_RATE_HZ;
                c.symbols_per_bit = 20;
                c.code_length_chips = static_cast<int>(GALILEO_E5A_CODE_LENGTH_CHIPS);
                c.secondary = p.track_pilot;  // the per-PRN E5a-Q secondary code is attached in start_tracking
                c.interchange_iq = p.track_pilot;
            }
        else if (p.system == 'C' && sig == "B1")
            {
                c.signal_carrier_freq = BEIDOU_B1I_FREQ_HZ;
                c.code_period = BEIDOU_B1I_CODE_PERIOD;
                c.code_chip_rate = BEIDOU_B1I_CODE_RATE_HZ;
                c.code_length_chips = static_cast<int>(BEIDOU_B1I_CODE_LENGTH_CHIPS);
                c.symbols_per_bit = BEIDOU_B1I_TELEMETRY_SYMBOLS_PER_BIT;
                c.secondary = true;
                c.track_pilot = false;
                c.secondary_code = BEIDOU_B1I_SECONDARY_CODE_STR;
            }
        else if (p.system == 'C' && sig == "B3")
            {
                c.signal_carrier_freq = BEIDOU_B3I_FREQ_HZ;
                c.code_period = BEIDOU_B3I_CODE_PERIOD;
                c.code_chip_rate = BEIDOU_B3I_CODE_RATE_HZ;
                c.code_length_chips = static_cast<int>(BEIDOU_B3I_CODE_LENGTH_CHIPS);
                c.symbols_per_bit = BEIDOU_B3I_TELEMETRY_SYMBOLS_PER_BIT;
                c.secondary = true;
                c.track_pilot = false;
                c.secondary_code = BEIDOU_B3I_SECONDARY_CODE_STR;
            }
    }

    gb200_service* s_;
    int stream_, item_type_;
    Dll_Pll_Conf trk_parameters;
    gb200::DllPllConf base_;
    gb200::B200_Multicorrelator_Real_Codes corr_;
    std::unique_ptr<gb200::DllPllVemlTracking<gb200::B200_Multicorrelator_Real_Codes>> loop_;
    Gnss_Synchro* d_acquisition_gnss_synchro = nullptr;
    uint32_t d_channel = 0;
    uint64_t d_sample_counter = 0;
    bool started_ = false, narrow_or_extended_ = false;
};

#endif
