// presentation.hpp -- headless analogue of the reference demo's front end (kernel.cu:19-139): the FPS overlay + out.png of
// display_image (30-43), the WASD camera translation it sketches in comments (52-110) and the mouse-look callback on_mouse
// (113-139).  There is no window and no OpenCV here, so
//   * key presses and mouse events come from a SCRIPT (one event per line) instead of cv::waitKey / cv::setMouseCallback;
//   * cv::imwrite becomes a self-contained PNG writer (8-bit RGB, zlib "stored" blocks -- any PNG reader opens it);
//   * cv::putText (Hershey vector font, an unpinned third-party rasteriser: SURVEY 8c) becomes a 5x7 bitmap font drawn at
//     the same anchor (10, 30), colour (0, 255, 0) BGR and roughly the same cap height -- same content, not the same pixels.
// The camera arithmetic is the reference's, operation for operation: a key moves the camera by apply_lre(invert_lre(pose), v)
// with |v| = 0.1 along a camera axis; a drag adds dx * 0.001 to yaw and dy * -0.001 to pitch (double arithmetic narrowed to
// float, as `pose->yaw += dx * 0.001` does).
#pragma once
#include "raytracer.hpp"
#include <cstdint>
#include <cstdio>
#include <deque>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

struct MouseParams                  // kernel.cu:19-28
{
    int last_x;
    int last_y;
    bool has_last = false;

    bool is_down = false;

    lre* pose;
};

namespace presentation {

enum { EVENT_MOUSEMOVE = 0, EVENT_LBUTTONDOWN = 1, EVENT_LBUTTONUP = 4 };      // the values of cv::MouseEventTypes

// ------------------------------------------------------------------ input: kernel.cu:52-110 (keys), 113-139 (mouse)
// Returns false when the key asks to quit ('q', kernel.cu:107-110: exit(0) there).
inline bool handle_key(int key, MouseParams& mouse_state)
{
    float3 step;
    if (key == 'w') step = make_float3(0, 0.1, 0);
    else if (key == 's') step = make_float3(0, -0.1, 0);
    else if (key == 'a') step = make_float3(-0.1, 0.0, 0);
    else if (key == 'd') step = make_float3(0.1, 0.0, 0);
    else return key != 'q';
    lre inv_camera_pose = invert_lre(*mouse_state.pose);
    float3 new_pos = apply_lre(inv_camera_pose, step);
    mouse_state.pose->x = new_pos.x;
    mouse_state.pose->y = new_pos.y;
    mouse_state.pose->z = new_pos.z;
    return true;
}

inline void on_mouse(int event, int x, int y, int, void* param)
{
    MouseParams* mouse_state = static_cast<MouseParams*>(param);
    if (event == EVENT_LBUTTONDOWN) mouse_state->is_down = true;
    else if (event == EVENT_LBUTTONUP) mouse_state->is_down = false;
    else if (event == EVENT_MOUSEMOVE) {
        if (mouse_state->has_last && mouse_state->is_down) {
            int dx = x - mouse_state->last_x;
            int dy = y - mouse_state->last_y;
            mouse_state->pose->yaw += dx * 0.001;
            mouse_state->pose->pitch += dy * -0.001;
        }
        mouse_state->last_x = x;
        mouse_state->last_y = y;
        mouse_state->has_last = true;
    }
}

// A scripted event source: the stand-in for the window's event queue.  Lines: `key <char>`, `down <x> <y>`, `up <x> <y>`,
// `move <x> <y>`, `frame` (nothing more happens during this frame's waitKey), `#` comments.
struct Script {
    struct Event { int kind; int a, b; };      // kind: 0 key(a), 1 mouse(event a at x,y packed in b), 2 end of frame
    std::deque<Event> events;
    std::vector<int> xs, ys;

    bool load(const std::string& path)
    {
        std::ifstream in(path);
        if (!in) return false;
        std::string line;
        while (std::getline(in, line)) {
            std::istringstream ss(line);
            std::string w;
            if (!(ss >> w) || w[0] == '#') continue;
            int x = 0, y = 0;
            if (w == "key") { std::string k; ss >> k; if (!k.empty()) events.push_back({ 0, (int)(unsigned char)k[0], 0 }); }
            else if (w == "frame") events.push_back({ 2, 0, 0 });
            else if (w == "down" || w == "up" || w == "move") {
                ss >> x >> y;
                xs.push_back(x); ys.push_back(y);
                events.push_back({ 1, w == "down" ? EVENT_LBUTTONDOWN : (w == "up" ? EVENT_LBUTTONUP : EVENT_MOUSEMOVE), (int)xs.size() - 1 });
            }
        }
        return true;
    }

    // One frame's worth of events: mouse events are delivered to on_mouse, the first key is handled like cv::waitKey(1)'s
    // return value (at most one key per frame).  Returns false on 'q'.
    bool pump(MouseParams& mouse_state)
    {
        while (!events.empty()) {
            Event e = events.front();
            events.pop_front();
            if (e.kind == 2) return true;
            if (e.kind == 1) on_mouse(e.a, xs[(size_t)e.b], ys[(size_t)e.b], 0, &mouse_state);
            else return handle_key(e.a, mouse_state);
        }
        return true;
    }
};

// ------------------------------------------------------------------ overlay: cv::putText stand-in
inline const unsigned char* glyph(char ch)
{
    static const struct { char c; unsigned char rows[7]; } G[] = {
        { '0', { 0x0E, 0x11, 0x13, 0x15, 0x19, 0x11, 0x0E } }, { '1', { 0x04, 0x0C, 0x04, 0x04, 0x04, 0x04, 0x0E } },
        { '2', { 0x0E, 0x11, 0x01, 0x02, 0x04, 0x08, 0x1F } }, { '3', { 0x1F, 0x02, 0x04, 0x02, 0x01, 0x11, 0x0E } },
        { '4', { 0x02, 0x06, 0x0A, 0x12, 0x1F, 0x02, 0x02 } }, { '5', { 0x1F, 0x10, 0x1E, 0x01, 0x01, 0x11, 0x0E } },
        { '6', { 0x06, 0x08, 0x10, 0x1E, 0x11, 0x11, 0x0E } }, { '7', { 0x1F, 0x01, 0x02, 0x04, 0x08, 0x08, 0x08 } },
        { '8', { 0x0E, 0x11, 0x11, 0x0E, 0x11, 0x11, 0x0E } }, { '9', { 0x0E, 0x11, 0x11, 0x0F, 0x01, 0x02, 0x0C } },
        { 'F', { 0x1F, 0x10, 0x10, 0x1E, 0x10, 0x10, 0x10 } }, { 'P', { 0x1E, 0x11, 0x11, 0x1E, 0x10, 0x10, 0x10 } },
        { 'S', { 0x0F, 0x10, 0x10, 0x0E, 0x01, 0x01, 0x1E } }, { ':', { 0x00, 0x0C, 0x0C, 0x00, 0x0C, 0x0C, 0x00 } },
        { '.', { 0x00, 0x00, 0x00, 0x00, 0x00, 0x0C, 0x0C } }, { '-', { 0x00, 0x00, 0x00, 0x1F, 0x00, 0x00, 0x00 } },
        { '+', { 0x00, 0x04, 0x04, 0x1F, 0x04, 0x04, 0x00 } }, { 'e', { 0x00, 0x00, 0x0E, 0x11, 0x1F, 0x10, 0x0E } },
        { 'i', { 0x04, 0x00, 0x0C, 0x04, 0x04, 0x04, 0x0E } }, { 'n', { 0x00, 0x00, 0x16, 0x19, 0x11, 0x11, 0x11 } },
        { 'f', { 0x06, 0x09, 0x08, 0x1C, 0x08, 0x08, 0x08 } }, { 'a', { 0x00, 0x00, 0x0E, 0x01, 0x0F, 0x11, 0x0F } },
        { ' ', { 0x00, 0x00, 0x00, 0x00, 0x00, 0x00, 0x00 } },
    };
    static const unsigned char box[7] = { 0x1F, 0x11, 0x11, 0x11, 0x11, 0x11, 0x1F };
    for (const auto& g : G) if (g.c == ch) return g.rows;
    return box;
}

// Draws `text` with its baseline-left corner at (x, y) -- cv::putText's `org` -- into a packed BGR image.
inline void put_text(unsigned char* bgr, int width, int height, const std::string& text, int x, int y, int scale,
                     unsigned char b, unsigned char g, unsigned char r)
{
    const int top = y - 7 * scale;
    for (size_t i = 0; i < text.size(); i++) {
        const unsigned char* rows = glyph(text[i]);
        const int gx = x + (int)i * 6 * scale;
        for (int ry = 0; ry < 7 * scale; ry++)
            for (int rx = 0; rx < 5 * scale; rx++) {
                if (!((rows[ry / scale] >> (4 - rx / scale)) & 1)) continue;
                const int px = gx + rx, py = top + ry;
                if (px < 0 || py < 0 || px >= width || py >= height) continue;
                unsigned char* p = bgr + ((size_t)py * width + px) * 3;
                p[0] = b; p[1] = g; p[2] = r;
            }
    }
}

// ------------------------------------------------------------------ cv::imwrite("out.png") stand-in
inline uint32_t crc32_update(uint32_t crc, const unsigned char* p, size_t n)
{
    static uint32_t table[256];
    static bool ready = false;
    if (!ready) {
        for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; }
        ready = true;
    }
    for (size_t i = 0; i < n; i++) crc = table[(crc ^ p[i]) & 0xFF] ^ (crc >> 8);
    return crc;
}

inline bool write_png(const std::string& path, const unsigned char* bgr, int width, int height)
{
    std::vector<unsigned char> raw((size_t)height * ((size_t)width * 3 + 1));
    for (int y = 0; y < height; y++) {
        unsigned char* row = &raw[(size_t)y * ((size_t)width * 3 + 1)];
        row[0] = 0;                                                        // filter: none
        const unsigned char* src = bgr + (size_t)y * width * 3;
        for (int x = 0; x < width; x++) { row[1 + 3 * x] = src[3 * x + 2]; row[2 + 3 * x] = src[3 * x + 1]; row[3 + 3 * x] = src[3 * x]; }
    }
    std::vector<unsigned char> z;                                          // zlib stream of stored (uncompressed) deflate blocks
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, bsum = 0;
    for (size_t off = 0; off < raw.size() || off == 0; off += 65535) {
        const size_t len = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + len >= raw.size() ? 1 : 0);
        z.push_back((unsigned char)(len & 0xFF)); z.push_back((unsigned char)(len >> 8));
        z.push_back((unsigned char)(~len & 0xFF)); z.push_back((unsigned char)((~len >> 8) & 0xFF));
        z.insert(z.end(), raw.begin() + (long)off, raw.begin() + (long)(off + len));
        for (size_t i = off; i < off + len; i++) { a = (a + raw[i]) % 65521u; bsum = (bsum + a) % 65521u; }
        if (raw.empty()) break;
    }
    const uint32_t adler = (bsum << 16) | a;
    for (int k = 3; k >= 0; k--) z.push_back((unsigned char)(adler >> (8 * k)));

    FILE* f = fopen(path.c_str(), "wb");
    if (!f) return false;
    auto chunk = [&](const char type[4], const unsigned char* data, size_t n) {
        unsigned char len[4] = { (unsigned char)(n >> 24), (unsigned char)(n >> 16), (unsigned char)(n >> 8), (unsigned char)n };
        fwrite(len, 1, 4, f);
        fwrite(type, 1, 4, f);
        if (n) fwrite(data, 1, n, f);
        uint32_t crc = crc32_update(0xFFFFFFFFu, (const unsigned char*)type, 4);
        crc = crc32_update(crc, data, n) ^ 0xFFFFFFFFu;
        unsigned char c[4] = { (unsigned char)(crc >> 24), (unsigned char)(crc >> 16), (unsigned char)(crc >> 8), (unsigned char)crc };
        fwrite(c, 1, 4, f);
    };
    static const unsigned char sig[8] = { 0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A };
    fwrite(sig, 1, 8, f);
    unsigned char ihdr[13] = { (unsigned char)(width >> 24), (unsigned char)(width >> 16), (unsigned char)(width >> 8), (unsigned char)width,
                               (unsigned char)(height >> 24), (unsigned char)(height >> 16), (unsigned char)(height >> 8), (unsigned char)height,
                               8, 2, 0, 0, 0 };
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", z.data(), z.size());
    chunk("IEND", nullptr, 0);
    return fclose(f) == 0;
}

// The host half of display_image (kernel.cu:39-43): overlay the frame rate, write the PNG.
inline bool present_host(unsigned char* img_cpu_bgr, int width, int height, double fps, const std::string& path = "out.png")
{
    std::string fps_text = "FPS: " + std::to_string(fps);
    put_text(img_cpu_bgr, width, height, fps_text, 10, 30, 3, 0, 255, 0);
    return write_png(path, img_cpu_bgr, width, height);
}

inline Script& script() { static Script s; return s; }                    // the demo's event source (empty = no input)
inline std::string& out_path() { static std::string p = "out.png"; return p; }

}  // namespace presentation

using presentation::on_mouse;

// kernel.cu:30-111.  Returns normally where the reference does; a scripted 'q' exits like the reference's (commented) one.
inline void display_image(uchar3* d_img, int width, int height, size_t pitch, double fps, MouseParams& mouse_state)
{
    std::vector<unsigned char> img_cpu((size_t)width * height * 3);       // GpuMat::download
    cudaMemcpy2D(img_cpu.data(), (size_t)width * 3, d_img, pitch, (size_t)width * 3, (size_t)height, cudaMemcpyDeviceToHost);
    if (!presentation::present_host(img_cpu.data(), width, height, fps, presentation::out_path()))
        std::cerr << "cannot write " << presentation::out_path() << std::endl;
    if (!presentation::script().pump(mouse_state)) std::exit(0);
}
