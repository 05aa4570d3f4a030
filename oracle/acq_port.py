"""TEST / BASELINE INFRASTRUCTURE ONLY -- numpy + scipy.fft restatement of pcps_acquisition::acquisition_core
used as the TIMED CPU baseline of bench.py ("kind": "port").  The C oracle (gnss_oracle.c) is the parity
checker; this file exists because a fair CPU time needs an FFTW-class FFT and FFTW itself is not installed:
scipy.fft (pocketfft, float32, workers = all cores) stands in -- "FFTW-class CPU FFT, not FFTW".

Follows src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc:
  set_local_code :239-274, update_local_carrier :296-310 (float32-accumulated phase like
  volk_gnsssdr_s32f_sincos_32fc_generic), acquisition_core :668-770, max_to_input_power_statistic :565-596.
"""
import numpy as np
import scipy.fft as sfft


class AcqPort:
    def __init__(self, fs, n_fft, consumed, doppler_max, doppler_step, workers=-1, group=4):
        self.fs, self.n, self.consumed = fs, n_fft, consumed
        self.group = group
        self.workers = workers
        self.num_bins = int(np.ceil(2 * doppler_max / doppler_step))
        dopp = -doppler_max + doppler_step * np.arange(self.num_bins)
        # init(): the reference stores one wipe-off table per bin (:353-359)
        steps = (6.283185307179586 * dopp.astype(np.float32).astype(np.float64) / np.float32(fs)).astype(np.float32)
        ph = np.cumsum(np.broadcast_to(-steps[:, None], (self.num_bins, n_fft)), axis=1, dtype=np.float32)
        ph = np.concatenate([np.zeros((self.num_bins, 1), np.float32), ph[:, :-1]], axis=1)
        self.wipe = (np.cos(ph) + 1j * np.sin(ph)).astype(np.complex64)
        self.doppler = dopp

    def set_local_code(self, code_c64):
        buf = np.zeros(self.n, np.complex64)
        buf[: len(code_c64)] = code_c64
        return np.conj(sfft.fft(buf, workers=self.workers)).astype(np.complex64)

    def run(self, samples_c64, code_ffts):
        """One dwell for every code in `code_ffts` (list) the way the reference does it: each channel repeats the
        wipe-off and forward FFT itself (one pcps_acquisition block per channel)."""
        x = np.zeros(self.n, np.complex64)
        x[: self.consumed] = samples_c64[: self.consumed]
        power = np.float32(np.sum(x.real * x.real + x.imag * x.imag, dtype=np.float32) / np.float32(self.n))
        out = []
        nf = np.float32(self.n) * np.float32(self.n)
        for cf in code_ffts:
            # the Doppler loop in cache-sized groups of bins (same arithmetic per bin; a whole-grid pass would stream 20 MB arrays
            # through DRAM five times): the first maximum over bins in ascending order wins, strict '>' (:576-583)
            best_mag, best_bin, best_idx = np.float32(0), 0, 0
            for b0 in range(0, self.num_bins, self.group):
                wiped = x[None, :] * self.wipe[b0: b0 + self.group]      # volk_32fc_x2_multiply_32fc per bin (:717)
                spec = sfft.fft(wiped, axis=1, workers=self.workers, overwrite_x=True)  # d_fft_if->execute() (:721)
                spec *= cf[None, :]                                      # (:724)
                y = sfft.ifft(spec, axis=1, workers=self.workers, norm="forward", overwrite_x=True)  # un-normalised backward (:727)
                mag = y.real * y.real + y.imag * y.imag                 # (:733)
                idx = np.argmax(mag, axis=1)                            # volk_gnsssdr_32f_index_max_32u per bin (:576)
                peak = mag[np.arange(mag.shape[0]), idx]
                k = int(np.argmax(peak))
                if peak[k] > best_mag:
                    best_mag, best_bin, best_idx = peak[k], b0 + k, int(idx[k])
            stat = np.float32(best_mag) / (nf * nf) / power
            out.append((best_idx, int(self.doppler[best_bin]), float(stat)))
        return out
