// presentation_test.cpp -- host-only check of presentation.hpp (no GPU needed): replays an input script against a camera
// pose, printing the pose after every frame, then overlays the frame rate on a synthetic BGR image and writes the PNG.
//   presentation_test <script.txt> <frames> <out.png> <width> <height> <fps>
#include "presentation.hpp"

int main(int argc, char** argv)
{
    if (argc < 7) { std::cerr << "usage: presentation_test <script> <frames> <out.png> <width> <height> <fps>" << std::endl; return 2; }
    lre pose;
    pose.x = -1; pose.y = -4; pose.z = 2;                                  // kernel.cu:168-170
    MouseParams mouse_state;
    mouse_state.pose = &pose;
    presentation::Script s;
    if (!s.load(argv[1])) { std::cerr << "cannot read " << argv[1] << std::endl; return 1; }
    const int frames = atoi(argv[2]);
    for (int f = 0; f < frames; f++) {
        const bool go_on = s.pump(mouse_state);
        printf("%d %.9g %.9g %.9g %.9g %.9g %.9g\n", f, pose.x, pose.y, pose.z, pose.yaw, pose.pitch, pose.roll);
        if (!go_on) { printf("quit\n"); break; }
    }
    const int w = atoi(argv[4]), h = atoi(argv[5]);
    std::vector<unsigned char> img((size_t)w * h * 3);
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) { unsigned char* p = &img[((size_t)y * w + x) * 3]; p[0] = (unsigned char)(x & 0xFF); p[1] = (unsigned char)(y & 0x7F); p[2] = (unsigned char)((x + y) & 0xFF); }
    return presentation::present_host(img.data(), w, h, atof(argv[6]), argv[3]) ? 0 : 1;
}
