// Minimal image file I/O for the headless sample programs (the reference's samples use cv::imread / cv::imshow, which
// need OpenCV and a display; neither exists here).  Reads binary PGM/PPM (P5/P6, maxval 255) and 8-bit non-interlaced
// PNG (gray, gray+alpha, RGB, RGBA; zlib for inflate); writes PGM/PPM and PNG.  Pixels come back in FILE channel order
// (RGB[A]); the device-side vip_convert_to_bgr_u8 turns them into the BGR layout cv::imread(IMREAD_COLOR) would give.
#ifndef VIP_APPS_IMAGE_IO_HPP
#define VIP_APPS_IMAGE_IO_HPP

#include <zlib.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace vipio {

struct Image {
    int width = 0, height = 0, channels = 0;  // channels: 1 gray, 2 gray+alpha, 3 RGB, 4 RGBA
    std::vector<std::uint8_t> data;           // tightly packed rows
    bool empty() const { return data.empty(); }
};

inline bool read_file(const std::string& path, std::vector<std::uint8_t>& out) {
    std::FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) return false;
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    out.resize(n > 0 ? static_cast<size_t>(n) : 0);
    const size_t got = out.empty() ? 0 : std::fread(out.data(), 1, out.size(), f);
    std::fclose(f);
    return got == out.size() && !out.empty();
}

// ------------------------------------------------------------------------------------------------ PNM
inline bool decode_pnm(const std::vector<std::uint8_t>& buf, Image& img) {
    if (buf.size() < 7 || buf[0] != 'P' || (buf[1] != '5' && buf[1] != '6')) return false;
    size_t pos = 2;
    int vals[3], nv = 0;
    while (nv < 3 && pos < buf.size()) {
        if (buf[pos] == '#') {
            while (pos < buf.size() && buf[pos] != '\n') ++pos;
        } else if (buf[pos] >= '0' && buf[pos] <= '9') {
            int v = 0;
            while (pos < buf.size() && buf[pos] >= '0' && buf[pos] <= '9') v = v * 10 + (buf[pos++] - '0');
            vals[nv++] = v;
            continue;
        }
        ++pos;
    }
    if (nv != 3 || vals[2] != 255 || vals[0] <= 0 || vals[1] <= 0) return false;
    ++pos;  // exactly one whitespace byte follows maxval
    img.width = vals[0];
    img.height = vals[1];
    img.channels = buf[1] == '5' ? 1 : 3;
    const size_t need = static_cast<size_t>(img.width) * img.height * img.channels;
    if (buf.size() - pos < need) return false;
    img.data.assign(buf.begin() + pos, buf.begin() + pos + need);
    return true;
}

// ------------------------------------------------------------------------------------------------ PNG
inline std::uint32_t be32(const std::uint8_t* p) {
    return (std::uint32_t(p[0]) << 24) | (std::uint32_t(p[1]) << 16) | (std::uint32_t(p[2]) << 8) | p[3];
}

inline int paeth(int a, int b, int c) {
    const int p = a + b - c, pa = std::abs(p - a), pb = std::abs(p - b), pc = std::abs(p - c);
    return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

inline bool decode_png(const std::vector<std::uint8_t>& buf, Image& img) {
    static const std::uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    if (buf.size() < 33 || std::memcmp(buf.data(), sig, 8) != 0) return false;
    size_t pos = 8;
    std::vector<std::uint8_t> idat;
    int color_type = -1;
    while (pos + 12 <= buf.size()) {
        const std::uint32_t len = be32(&buf[pos]);
        const char* type = reinterpret_cast<const char*>(&buf[pos + 4]);
        if (pos + 12 + len > buf.size()) return false;
        const std::uint8_t* d = &buf[pos + 8];
        if (!std::memcmp(type, "IHDR", 4)) {
            if (len < 13) return false;
            img.width = static_cast<int>(be32(d));
            img.height = static_cast<int>(be32(d + 4));
            color_type = d[9];
            if (d[8] != 8 || d[10] != 0 || d[11] != 0 || d[12] != 0) return false;  // 8-bit, non-interlaced only
        } else if (!std::memcmp(type, "IDAT", 4)) {
            idat.insert(idat.end(), d, d + len);
        } else if (!std::memcmp(type, "IEND", 4)) {
            break;
        }
        pos += 12 + len;
    }
    switch (color_type) {
        case 0: img.channels = 1; break;
        case 2: img.channels = 3; break;
        case 4: img.channels = 2; break;
        case 6: img.channels = 4; break;
        default: return false;  // palette images are not supported
    }
    if (img.width <= 0 || img.height <= 0) return false;
    const size_t bpp = img.channels, stride = static_cast<size_t>(img.width) * bpp;
    std::vector<std::uint8_t> raw((stride + 1) * img.height);
    uLongf raw_len = static_cast<uLongf>(raw.size());
    if (uncompress(raw.data(), &raw_len, idat.data(), static_cast<uLong>(idat.size())) != Z_OK || raw_len != raw.size())
        return false;
    img.data.assign(stride * img.height, 0);
    for (int y = 0; y < img.height; ++y) {
        const std::uint8_t* in = &raw[(stride + 1) * y];
        std::uint8_t* cur = &img.data[stride * y];
        const std::uint8_t* up = y ? cur - stride : nullptr;
        const int filter = in[0];
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? cur[i - bpp] : 0, b = up ? up[i] : 0, c = (up && i >= bpp) ? up[i - bpp] : 0;
            int v = in[1 + i];
            switch (filter) {
                case 0: break;
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) / 2; break;
                case 4: v += paeth(a, b, c); break;
                default: return false;
            }
            cur[i] = static_cast<std::uint8_t>(v);
        }
    }
    return true;
}

inline Image imread(const std::string& path) {
    Image img;
    std::vector<std::uint8_t> buf;
    if (!read_file(path, buf)) return img;
    if (!(decode_png(buf, img) || decode_pnm(buf, img))) img = Image();
    return img;
}

// ------------------------------------------------------------------------------------------------ writers
inline bool write_pnm(const std::string& path, const std::uint8_t* data, int width, int height, int channels) {
    if (channels != 1 && channels != 3) return false;
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    std::fprintf(f, "P%c\n%d %d\n255\n", channels == 1 ? '5' : '6', width, height);
    const size_t n = static_cast<size_t>(width) * height * channels;
    const bool ok = std::fwrite(data, 1, n, f) == n;
    std::fclose(f);
    return ok;
}

inline void put_chunk(std::vector<std::uint8_t>& out, const char* type, const std::uint8_t* d, size_t len) {
    const std::uint8_t l[4] = {std::uint8_t(len >> 24), std::uint8_t(len >> 16), std::uint8_t(len >> 8), std::uint8_t(len)};
    out.insert(out.end(), l, l + 4);
    const size_t start = out.size();
    out.insert(out.end(), type, type + 4);
    if (len) out.insert(out.end(), d, d + len);
    const std::uint32_t crc = static_cast<std::uint32_t>(crc32(0L, &out[start], static_cast<uInt>(4 + len)));
    const std::uint8_t c[4] = {std::uint8_t(crc >> 24), std::uint8_t(crc >> 16), std::uint8_t(crc >> 8), std::uint8_t(crc)};
    out.insert(out.end(), c, c + 4);
}

inline bool write_png(const std::string& path, const std::uint8_t* data, int width, int height, int channels) {
    if (channels != 1 && channels != 3) return false;
    const size_t stride = static_cast<size_t>(width) * channels;
    std::vector<std::uint8_t> raw((stride + 1) * height);
    for (int y = 0; y < height; ++y) {
        raw[(stride + 1) * y] = 0;
        std::memcpy(&raw[(stride + 1) * y + 1], data + stride * y, stride);
    }
    uLongf clen = compressBound(static_cast<uLong>(raw.size()));
    std::vector<std::uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), static_cast<uLong>(raw.size()), 6) != Z_OK) return false;
    std::vector<std::uint8_t> out = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::uint8_t ihdr[13] = {std::uint8_t(width >> 24), std::uint8_t(width >> 16), std::uint8_t(width >> 8), std::uint8_t(width),
                             std::uint8_t(height >> 24), std::uint8_t(height >> 16), std::uint8_t(height >> 8), std::uint8_t(height),
                             8, std::uint8_t(channels == 1 ? 0 : 2), 0, 0, 0};
    put_chunk(out, "IHDR", ihdr, 13);
    put_chunk(out, "IDAT", comp.data(), clen);
    put_chunk(out, "IEND", nullptr, 0);
    std::FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) return false;
    const bool ok = std::fwrite(out.data(), 1, out.size(), f) == out.size();
    std::fclose(f);
    return ok;
}

// `rgb` is in file order (RGB); extension decides the format (.png, else PNM)
inline bool imwrite(const std::string& path, const std::uint8_t* rgb, int width, int height, int channels) {
    const bool png = path.size() >= 4 && path.compare(path.size() - 4, 4, ".png") == 0;
    return png ? write_png(path, rgb, width, height, channels) : write_pnm(path, rgb, width, height, channels);
}

}  // namespace vipio
#endif
