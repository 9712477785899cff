// Minimal grayscale image IO for the command-line tools: binary/ASCII PGM in (the reference reads through
// cv::imread(..., IMREAD_GRAYSCALE), tool/detection_example.cc:25; OpenCV's C++ package is not in this image),
// PGM / PPM out (instead of cv::imshow, which needs a display). Values stay 0..255 floats, like
// OpencvUtils::cvMatToImage<float> of an 8-bit Mat (cvUtils/ConversionImpl.hpp:25-27).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

#include "sift_cuda/types/HostImage.hh"

namespace cli {

inline bool read_pgm(const std::string& path, Imagef& out) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::string magic;
    f >> magic;
    if (magic != "P5" && magic != "P2") return false;
    auto next_int = [&f]() {
        for (;;) {
            f >> std::ws;
            if (f.peek() != '#') break;
            std::string line;
            std::getline(f, line);
        }
        int v = -1;
        f >> v;
        return v;
    };
    const int w = next_int(), h = next_int(), maxv = next_int();
    if (w <= 0 || h <= 0 || maxv <= 0 || maxv > 65535) return false;
    out = Imagef(h, w);
    std::vector<float>& d = *out.m_data;
    if (magic == "P2") {
        for (size_t i = 0; i < d.size(); i++) d[i] = float(next_int());
    } else {
        f.get();  // the single whitespace after maxval
        const size_t bpp = maxv > 255 ? 2 : 1;
        std::vector<unsigned char> raw(d.size() * bpp);
        f.read(reinterpret_cast<char*>(raw.data()), std::streamsize(raw.size()));
        if (size_t(f.gcount()) != raw.size()) return false;
        for (size_t i = 0; i < d.size(); i++) d[i] = bpp == 1 ? float(raw[i]) : float((raw[2 * i] << 8) | raw[2 * i + 1]);
    }
    if (maxv != 255)  // bring other depths onto the 0..255 range the thresholds assume
        for (float& v : d) v = v * 255.f / float(maxv);
    return true;
}

inline bool write_pgm(const std::string& path, const Imagef& img) {
    const int h = int(img.m_image_size.row), w = int(img.m_image_size.col);
    std::ofstream f(path, std::ios::binary);
    if (!f) return false;
    f << "P5\n" << w << " " << h << "\n255\n";
    std::vector<unsigned char> raw(size_t(w) * h);
    for (size_t i = 0; i < raw.size(); i++) raw[i] = (unsigned char)std::min(255.f, std::max(0.f, (*img.m_data)[i]));
    f.write(reinterpret_cast<const char*>(raw.data()), std::streamsize(raw.size()));
    return bool(f);
}

// RGB canvas with the two frames side by side and one line per match (the job cv::drawMatches does in
// tool/extract_and_match_example.cc:94-95), written as binary PPM.
struct Canvas {
    int w = 0, h = 0;
    std::vector<unsigned char> rgb;
    Canvas(const Imagef& a, const Imagef& b) {
        const int ha = int(a.m_image_size.row), wa = int(a.m_image_size.col), hb = int(b.m_image_size.row), wb = int(b.m_image_size.col);
        w = wa + wb;
        h = std::max(ha, hb);
        rgb.assign(size_t(w) * h * 3, 0);
        auto blit = [this](const Imagef& im, int x0, int iw, int ih) {
            for (int y = 0; y < ih; y++)
                for (int x = 0; x < iw; x++) {
                    const unsigned char v = (unsigned char)std::min(255.f, std::max(0.f, (*im.m_data)[size_t(y) * iw + x]));
                    unsigned char* p = &rgb[(size_t(y) * w + x0 + x) * 3];
                    p[0] = p[1] = p[2] = v;
                }
        };
        blit(a, 0, wa, ha);
        blit(b, wa, wb, hb);
    }
    void line(float x0, float y0, float x1, float y1, unsigned char r, unsigned char g, unsigned char b) {
        const int n = int(std::max(std::fabs(x1 - x0), std::fabs(y1 - y0))) + 1;
        for (int i = 0; i <= n; i++) {
            const int x = int(std::lround(x0 + (x1 - x0) * i / n)), y = int(std::lround(y0 + (y1 - y0) * i / n));
            if (x < 0 || y < 0 || x >= w || y >= h) continue;
            unsigned char* p = &rgb[(size_t(y) * w + x) * 3];
            p[0] = r; p[1] = g; p[2] = b;
        }
    }
    bool write_ppm(const std::string& path) const {
        std::ofstream f(path, std::ios::binary);
        if (!f) return false;
        f << "P6\n" << w << " " << h << "\n255\n";
        f.write(reinterpret_cast<const char*>(rgb.data()), std::streamsize(rgb.size()));
        return bool(f);
    }
};

// "-p value" / "--path value" / "--path=value" option lookup (the reference uses CLI11, which is not in this image).
inline std::string option(int argc, char** argv, const std::string& short_name, const std::string& long_name, const std::string& def = "") {
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        if ((a == short_name || a == long_name) && i + 1 < argc) return argv[i + 1];
        if (a.rfind(long_name + "=", 0) == 0) return a.substr(long_name.size() + 1);
    }
    return def;
}
inline bool flag(int argc, char** argv, const std::string& short_name, const std::string& long_name) {
    for (int i = 1; i < argc; i++)
        if (short_name == argv[i] || long_name == argv[i]) return true;
    return false;
}

}  // namespace cli
