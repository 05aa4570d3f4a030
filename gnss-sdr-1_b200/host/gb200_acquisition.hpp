// B200_Pcps_Acquisition: host mirror of the pcps_acquisition block's control surface
// (src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.h:93-250 and the AcquisitionInterface calls the Channel
// makes on its adapter, src/core/interfaces/acquisition_interface.h:56-72) on top of the C ABI.  GNU Radio's scheduler is
// replaced by an explicit run(samp_count): the block's own buffering state machine (general_work, :941-1054) only decides
// WHICH d_consumed_samples items to search, and here those are named by their absolute sample index in the device ring.
#pragma once
#include <cmath>
#include <complex>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/gnss_b200.h"
#include "gb200_matfile.hpp"

namespace gb200
{
// the Gnss_Synchro fields acquisition writes (gnss_synchro.h:54-58)
struct AcqSynchro
{
    double Acq_delay_samples = 0.0;
    double Acq_doppler_hz = 0.0;
    uint64_t Acq_samplestamp_samples = 0;
    uint32_t Acq_doppler_step = 0;
    bool Flag_valid_acquisition = false;
    // identity of the satellite searched (gnss_synchro.h:44-48); only the dump file name and its PRN variable read them here
    char System = 'G';
    char Signal[3] = {'1', 'C', 0};
    uint32_t PRN = 0;
};

// = the adapters' calculate_threshold(pfa) (gps_l1_ca_pcps_acquisition.cc:262-279, galileo_e1_pcps_ambiguous_acquisition.cc:299-317):
// the (1 - pfa)^(1 / ncells) quantile of an exponential distribution with rate lambda = vector_length, which is
// -log1p(-p) / lambda (what boost::math::quantile evaluates).  The bin count loop is inclusive (<= doppler_max), unlike the
// block's own grid (pcps_acquisition.cc:326), and ncells is an unsigned 32-bit product, both as written there.
inline float pcps_threshold_from_pfa(float pfa, unsigned doppler_max, unsigned doppler_step, unsigned vector_length)
{
    unsigned frequency_bins = 0;
    if (doppler_step == 0) return 0.0f;
    for (int doppler = -static_cast<int>(doppler_max); doppler <= static_cast<int>(doppler_max); doppler += static_cast<int>(doppler_step)) frequency_bins++;
    const unsigned ncells = vector_length * frequency_bins;
    const double exponent = 1 / static_cast<double>(ncells);
    const double val = std::pow(1.0 - pfa, exponent);
    const double lambda = static_cast<double>(vector_length);
    return static_cast<float>(-std::log1p(-val) / lambda);
}

// Integer decimation of the acquisition resampler (gnss_flowgraph.cc:431-437; the adapters repeat it, e.g.
// gps_l1_ca_pcps_acquisition.cc:97-104): the largest divisor of fs_in not above fs_in / opt_acq_fs; 1 = no resampler.
inline uint32_t acq_resampler_decimation(int64_t fs_in, uint32_t opt_acq_fs)
{
    if (fs_in <= static_cast<int64_t>(opt_acq_fs)) return 1;
    uint32_t decimation = static_cast<uint32_t>(fs_in / opt_acq_fs);
    while (fs_in % decimation > 0) decimation--;
    return decimation;
}

// The flowgraph's resampler for one (signal, RF channel): taps = firdes::low_pass(1.0, fs, acq_fs / 2.1, acq_fs / 10, WIN_HAMMING)
// with acq_fs = fs / decimation (gnss_flowgraph.cc:437-447), attached to the source stream.  Returns the filter latency the
// flowgraph hands to set_resampler_latency(), (taps.size() - 1) / 2 (:476), or -1.
inline int attach_acq_resampler(gb200_service* s, int src_stream, int dst_stream, int64_t fs_in, uint32_t decimation)
{
    const double acq_fs = static_cast<double>(fs_in) / static_cast<double>(decimation);
    uint32_t n = 0;
    std::vector<float> taps(1);
    if (gb200_firdes_low_pass(1.0, static_cast<double>(fs_in), acq_fs / 2.1, acq_fs / 10, taps.data(), &n) != GB200_E_RANGE) return -1;
    taps.resize(n);
    if (gb200_firdes_low_pass(1.0, static_cast<double>(fs_in), acq_fs / 2.1, acq_fs / 10, taps.data(), &n) != GB200_OK) return -1;
    if (gb200_stream_set_decimator(s, src_stream, dst_stream, decimation, taps.data(), n) != GB200_OK) return -1;
    return static_cast<int>((taps.size() - 1) / 2);
}

class B200_Pcps_Acquisition
{
public:
    // conf = the Acq_Conf the adapter fills (acq_conf.h:39-68); item_type = GB200_ITEM_* of the stream searched
    B200_Pcps_Acquisition(gb200_service* service, const gb200_acq_cfg& conf, int stream_id) : s_(service), conf_(conf), stream_(stream_id)
    {
        ok_ = gb200_acq_create(s_, &conf_, &h_) == GB200_OK;
    }
    ~B200_Pcps_Acquisition()
    {
        if (h_) gb200_acq_destroy(h_);
    }
    B200_Pcps_Acquisition(const B200_Pcps_Acquisition&) = delete;
    B200_Pcps_Acquisition& operator=(const B200_Pcps_Acquisition&) = delete;
    bool ok() const { return ok_; }

    // ---- AcquisitionInterface, in the order Channel calls it (channel.cc:60-104) ----
    void set_channel(unsigned channel) { channel_ = channel; }
    void set_gnss_synchro(AcqSynchro* synchro) { synchro_ = synchro; }          // borrowed, like Gnss_Synchro* (pcps_acquisition.h:133)
    void set_doppler_max(unsigned doppler_max)
    {
        conf_.doppler_max = doppler_max;
        gb200_acq_set_doppler(h_, conf_.doppler_max, conf_.doppler_step, conf_.freq_offset_hz);
    }
    void set_doppler_step(unsigned doppler_step)
    {
        conf_.doppler_step = doppler_step;
        gb200_acq_set_doppler(h_, conf_.doppler_max, conf_.doppler_step, conf_.freq_offset_hz);
    }
    void set_threshold(float threshold)
    {
        conf_.threshold = threshold;
        gb200_acq_set_threshold(h_, threshold);
    }
    // dump / dump_filename / dump_channel of Acq_Conf (pcps_acquisition.cc:160-194): one .mat per finished search of `dump_channel`
    void set_dump(const std::string& dump_filename, unsigned dump_channel)
    {
        dump_filename_ = dump_filename;
        dump_channel_ = dump_channel;
        dump_ = !dump_filename.empty();
        gb200_acq_keep_grid(h_, dump_ ? 1 : 0);
    }
    const std::string& last_dump_file() const { return last_dump_file_; }
    void init() { gb200_acq_set_doppler(h_, conf_.doppler_max, conf_.doppler_step, conf_.freq_offset_hz); }  // (:313-370): grid rebuilt, nothing stored
    // = pcps_acquisition::set_local_code (:239-274): `code` has d_consumed_samples (or one code period, zero-padded inside) items
    bool set_local_code(const std::complex<float>* code, size_t n_items)
    {
        return gb200_acq_set_local_code(h_, reinterpret_cast<const float*>(code), n_items) == GB200_OK;
    }
    void set_state(int state)  // (:393-415): entering state 1 re-arms the search
    {
        state_ = state;
        if (state == 1) gb200_acq_reset(h_);
    }
    // Acq_Conf::use_automatic_resampler / resampler_ratio (the adapter fills them, gps_l1_ca_pcps_acquisition.cc:89-112) and
    // set_resampler_latency (pcps_acquisition.cc:232-236).  `conf` then describes the RESAMPLED stream (fs_in = resampled_fs,
    // which is also what update_local_carrier uses, :299-302) and `stream_id` names the decimator's destination stream.
    void set_automatic_resampler(float resampler_ratio)
    {
        use_automatic_resampler_ = true;
        resampler_ratio_ = resampler_ratio;
    }
    void set_resampler_latency(uint32_t latency_samples) { resampler_latency_samples_ = latency_samples; }
    void reset() { set_state(1); }  // ChannelFsm::start_acquisition -> adapter reset() -> set_active(true) (:226-229)
    void stop_acquisition() { state_ = 0; }
    signed int mag() const { return static_cast<signed int>(mag_); }
    unsigned fft_size() const { return gb200_acq_fft_size(h_); }
    unsigned consumed_samples() const { return gb200_acq_consumed_samples(h_); }

    // = acquisition_core(samp_count) (:668-929) on the items [samp_count, samp_count + d_consumed_samples) of the stream.
    // Returns the event the block would send: 1 positive, 2 negative (pcps_acquisition.cc:418-459), 0 none yet
    // (more dwells wanted, or step two armed: call run() again with the next / the same block).
    int run(uint64_t samp_count, gb200_acq_result* detail = nullptr)
    {
        if (state_ == 0) return 0;
        gb200_acq_result r;
        if (gb200_acq_run(h_, stream_, samp_count, &r) != GB200_OK) return -1;
        if (detail) *detail = r;
        mag_ = r.peak_mag;
        test_statistics_ = r.test_statistic;
        if (synchro_)
            {
                synchro_->Acq_delay_samples = r.acq_delay_samples;
                synchro_->Acq_doppler_hz = r.acq_doppler_hz;
                synchro_->Acq_samplestamp_samples = r.acq_samplestamp;
                synchro_->Acq_doppler_step = conf_.doppler_step;
                if (use_automatic_resampler_)
                    {
                        // (:756-762, :814-820) back to the rate the tracking blocks run at
                        synchro_->Acq_delay_samples = static_cast<double>(std::fmod(static_cast<float>(r.code_phase_index), conf_.samples_per_code)) * resampler_ratio_;
                        synchro_->Acq_delay_samples -= static_cast<double>(resampler_latency_samples_);
                        synchro_->Acq_samplestamp_samples = rint(static_cast<double>(r.acq_samplestamp) * resampler_ratio_);
                    }
            }
        if (r.decision != 3) num_dwells_++;  // d_num_noncoherent_integrations_counter (:711); the step-two pass does not count
        if (r.decision == 1 || r.decision == 0)
            {
                if (dump_ && channel_ == dump_channel_) dump_results(r);  // (:913-918)
                num_dwells_ = 0;
                state_ = 0;  // d_state = 0; d_active = false
                if (synchro_) synchro_->Flag_valid_acquisition = (r.decision == 1);
                return r.decision == 1 ? 1 : 2;
            }
        return 0;
    }
    float test_statistics() const { return test_statistics_; }
    std::string last_error() const { return gb200_last_error(s_); }

private:
    // = dump_results (pcps_acquisition.cc:462-562): same file name pattern, variable names, classes and dimensions; written as a
    // Level-5 MAT-file (see gb200_matfile.hpp).  The step-two narrow grid is not kept on the device and is not dumped.
    void dump_results(const gb200_acq_result& r)
    {
        dump_number_++;
        const AcqSynchro blank;
        const AcqSynchro& syn = synchro_ ? *synchro_ : blank;
        std::string filename = dump_filename_;
        filename += "_";
        filename += syn.System;
        filename += "_";
        filename += syn.Signal[0];
        filename += syn.Signal[1];
        filename += "_ch_" + std::to_string(channel_) + "_" + std::to_string(dump_number_) + "_sat_" + std::to_string(syn.PRN) + ".mat";
        const uint32_t bins = gb200_acq_num_doppler_bins(h_);
        const uint32_t eff = conf_.bit_transition_flag ? gb200_acq_fft_size(h_) / 2 : gb200_acq_fft_size(h_);
        std::vector<float> grid(static_cast<size_t>(bins) * eff);
        if (gb200_acq_read_grid(h_, grid.data()) != GB200_OK) return;
        MatFileWriter m(filename);
        if (!m.ok()) return;
        // the device grid is [bin][index] row-major, i.e. exactly the column-major (index, bin) matrix grid_ is (:743)
        m.write("acq_grid", MatFileWriter::kSingle, grid.data(), eff, bins);
        m.write_scalar<uint32_t>("doppler_max", MatFileWriter::kUint32, conf_.doppler_max);
        m.write_scalar<uint32_t>("doppler_step", MatFileWriter::kUint32, conf_.doppler_step);
        m.write_scalar<int32_t>("d_positive_acq", MatFileWriter::kInt32, r.decision == 1 ? 1 : 0);
        m.write_scalar<float>("acq_doppler_hz", MatFileWriter::kSingle, static_cast<float>(r.acq_doppler_hz));
        m.write_scalar<float>("acq_delay_samples", MatFileWriter::kSingle, static_cast<float>(r.acq_delay_samples));
        m.write_scalar<float>("test_statistic", MatFileWriter::kSingle, r.test_statistic);
        m.write_scalar<float>("threshold", MatFileWriter::kSingle, conf_.threshold);
        m.write_scalar<float>("input_power", MatFileWriter::kSingle, r.input_power);
        m.write_scalar<uint64_t>("sample_counter", MatFileWriter::kUint64, r.acq_samplestamp);
        m.write_scalar<uint32_t>("PRN", MatFileWriter::kUint32, syn.PRN);
        m.write_scalar<uint32_t>("num_dwells", MatFileWriter::kUint32, num_dwells_);
        m.close();
        last_dump_file_ = filename;
    }

    gb200_service* s_;
    gb200_acq_cfg conf_;
    gb200_acq* h_ = nullptr;
    AcqSynchro* synchro_ = nullptr;
    int stream_;
    unsigned channel_ = 0;
    int state_ = 0;
    bool use_automatic_resampler_ = false;
    float resampler_ratio_ = 1.0f;
    uint32_t resampler_latency_samples_ = 0;
    bool ok_ = false;
    float mag_ = 0.0f, test_statistics_ = 0.0f;
    bool dump_ = false;
    unsigned dump_channel_ = 0;
    uint32_t num_dwells_ = 0;
    long long dump_number_ = 0;
    std::string dump_filename_, last_dump_file_;
};
}  // namespace gb200
