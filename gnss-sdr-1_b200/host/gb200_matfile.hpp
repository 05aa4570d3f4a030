// Minimal MAT-file writer (Level 5, uncompressed numeric arrays) for the acquisition dump.
//
// The reference writes its acquisition dump through matio as a v7.3 (HDF5) file (pcps_acquisition.cc:479, Mat_CreateVer(...,
// MAT_FT_MAT73)); neither matio nor HDF5 is available here.  Its readers -- acquisition_dump_reader.cc (Mat_Open + Mat_VarRead by
// variable NAME, :40-126) and MATLAB's load() -- open Level-5 files the same way, so compatibility is by variable names, classes
// and dimensions, which this writer reproduces.  Column-major data, little-endian.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace gb200
{
class MatFileWriter
{
public:
    enum Class : uint8_t
    {
        kSingle = 7,
        kInt32 = 12,
        kUint32 = 13,
        kUint64 = 15
    };

    explicit MatFileWriter(const std::string& filename) : f_(std::fopen(filename.c_str(), "wb"))
    {
        if (!f_) return;
        char header[128];
        std::memset(header, ' ', sizeof header);
        const char text[] = "MATLAB 5.0 MAT-file, Platform: B200, Created by: gnss-sdr-1_b200";
        std::memcpy(header, text, sizeof text - 1);
        std::memset(header + 116, 0, 8);  // subsystem data offset
        const uint16_t version = 0x0100;
        std::memcpy(header + 124, &version, 2);
        header[126] = 'I';  // little-endian marker as a reader on a little-endian machine sees it: "IM"
        header[127] = 'M';
        std::fwrite(header, 1, sizeof header, f_);
    }
    ~MatFileWriter() { close(); }
    MatFileWriter(const MatFileWriter&) = delete;
    MatFileWriter& operator=(const MatFileWriter&) = delete;
    bool ok() const { return f_ != nullptr; }
    void close()
    {
        if (f_) std::fclose(f_);
        f_ = nullptr;
    }

    // rows x cols real array of `cls`, `data` column-major
    bool write(const char* name, Class cls, const void* data, size_t rows, size_t cols)
    {
        if (!f_) return false;
        const uint32_t mi_type = cls == kSingle ? 7u : cls == kInt32 ? 5u : cls == kUint32 ? 6u : 13u;  // miSINGLE, miINT32, miUINT32, miUINT64
        const size_t elem = cls == kUint64 ? 8 : 4;
        const size_t data_bytes = rows * cols * elem;
        const size_t name_len = std::strlen(name);
        const size_t body = 16 + 16 + (8 + pad8(name_len)) + (8 + pad8(data_bytes));
        if (body > 0xFFFFFFFFull) return false;
        std::vector<unsigned char> buf;
        buf.reserve(8 + body - pad8(data_bytes));
        tag(buf, 14, static_cast<uint32_t>(body));  // miMATRIX
        tag(buf, 6, 8);                             // array flags: miUINT32 x 2
        u32(buf, cls);
        u32(buf, 0);
        tag(buf, 5, 8);  // dimensions: miINT32 x 2
        u32(buf, static_cast<uint32_t>(rows));
        u32(buf, static_cast<uint32_t>(cols));
        tag(buf, 1, static_cast<uint32_t>(name_len));  // name: miINT8
        buf.insert(buf.end(), name, name + name_len);
        buf.resize(buf.size() + pad8(name_len) - name_len, 0);
        tag(buf, mi_type, static_cast<uint32_t>(data_bytes));
        if (std::fwrite(buf.data(), 1, buf.size(), f_) != buf.size()) return false;
        if (data_bytes && std::fwrite(data, 1, data_bytes, f_) != data_bytes) return false;
        static const unsigned char zeros[8] = {};
        const size_t padn = pad8(data_bytes) - data_bytes;
        return padn == 0 || std::fwrite(zeros, 1, padn, f_) == padn;
    }
    template <class T>
    bool write_scalar(const char* name, Class cls, T v)
    {
        return write(name, cls, &v, 1, 1);
    }

private:
    static size_t pad8(size_t n) { return (n + 7) & ~static_cast<size_t>(7); }
    static void u32(std::vector<unsigned char>& b, uint32_t v)
    {
        unsigned char t[4];
        std::memcpy(t, &v, 4);
        b.insert(b.end(), t, t + 4);
    }
    static void tag(std::vector<unsigned char>& b, uint32_t type, uint32_t nbytes)
    {
        u32(b, type);
        u32(b, nbytes);
    }
    std::FILE* f_;
};
}  // namespace gb200
