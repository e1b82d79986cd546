// demo.cpp -- headless analogue of the reference's demo application (kernel.cu:141-302), written against the
// drop-in headers exactly the way an application written against AFIDclan/cuda-raytracing is: same includes, same
// calls, same frame loop (two render_scene launches, one cudaDeviceSynchronize, FPS print).  OpenCV's imwrite/imshow
// (kernel.cu:30-49) become a binary PPM (headless presentation); the git-ignored calibration_* assets
// become any OBJ (and optional PPM texture) given on the command line.
//
// When the output name ends in .png the frame loop presents like the reference's: display_image after every frame (FPS overlay,
// PNG rewritten each frame, kernel.cu:30-43,288) with keys and mouse events replayed from a script (presentation.hpp).
//
//   demo <mesh.obj> [out.ppm|out.png] [frames] [mode 0..3] [width height] [texture.ppm|-] [input-script.txt]
#include "Camera.h"
#include "Scene.h"
#include "OBJLoader.hpp"
#include "presentation.hpp"
#include <chrono>

static void write_ppm(const char* path, const uchar3* d_img, int width, int height, size_t pitch)
{
    std::vector<unsigned char> bgr((size_t)width * height * 3);
    cudaMemcpy2D(bgr.data(), (size_t)width * 3, d_img, pitch, (size_t)width * 3, (size_t)height, cudaMemcpyDeviceToHost);
    for (size_t i = 0; i + 2 < bgr.size(); i += 3) std::swap(bgr[i], bgr[i + 2]);        // BGR (CV_8UC3) -> RGB
    FILE* f = fopen(path, "wb");
    if (!f) { std::cerr << "cannot write " << path << std::endl; return; }
    fprintf(f, "P6\n%d %d\n255\n", width, height);
    fwrite(bgr.data(), 1, bgr.size(), f);
    fclose(f);
}

int main(int argc, char** argv)
{
    if (argc < 2) { std::cerr << "usage: demo <mesh.obj> [out.ppm|out.png] [frames] [mode] [width height] [texture.ppm|-] [input-script.txt]" << std::endl; return 2; }
    const char* out = argc > 2 ? argv[2] : "out.ppm";
    const std::string out_s(out);
    const bool present = out_s.size() > 4 && out_s.compare(out_s.size() - 4, 4, ".png") == 0;
    if (present) presentation::out_path() = out_s;
    if (argc > 8 && !presentation::script().load(argv[8])) { std::cerr << "cannot read " << argv[8] << std::endl; return 1; }
    const int frames = argc > 3 ? atoi(argv[3]) : 100;
    const int mode = argc > 4 ? atoi(argv[4]) : 0;
    int width = 1920, height = 1080;                                    // kernel.cu:155-156
    if (argc > 6) { width = atoi(argv[5]); height = atoi(argv[6]); }
    const float s = (float)width / 1920.0f;

    float4 D = make_float4(0.016233999489849514, -0.013875757716177956, 0.03264329940126211, -0.019561619947134234);   // kernel.cu:158
    float3x3 K = { { { 862.097835972576f * s, 0.0f, 998.1702383680802f * s }, { 0.0f, 862.1368447300727f * s, 569.6759403225842f * s }, { 0.0f, 0.0f, 1.0f } } };

    Camera camera = Camera(width, height, K, D);
    camera.pose.x = -1; camera.pose.y = -4; camera.pose.z = 2;          // kernel.cu:168-170
    camera.mode = mode; camera.max_depth = mode == 2 ? 4 : (mode == 3 ? 3 : 0); camera.spp = mode == 3 ? 4 : 1;

    Scene scene = Scene();
    Material glossy_red = Material();
    glossy_red.albedo = make_float3(0.1, 0.2, 0.9);
    glossy_red.roughness = 0.01;
    glossy_red.metallic = mode == 2 ? 0.5f : 0.0f;
    scene.add_material(glossy_red);
    Material matte = Material();
    matte.albedo = make_float3(0.9, 0.9, 0.9);
    matte.roughness = 0.3;
    if (argc > 7 && std::string(argv[7]) != "-") matte.upload_texture(std::string(argv[7]));
    scene.add_material(matte);

    MeshPrimitive mesh = OBJLoader::load(argv[1]);
    scene.add_mesh(mesh);

    MeshInstance a = MeshInstance(0, 0);
    scene.add_mesh_instance(a);
    MeshInstance b = MeshInstance(0, 1);
    b.pose.x = -2.6; b.pose.y = 1.48; b.pose.z = 0.73; b.pose.yaw = 0.6;
    scene.add_mesh_instance(b);

    scene.upload_to_device();
    mesh.bvh_top.print_stats();

    uchar3* d_img; size_t pitch;
    cudaMallocPitch(&d_img, &pitch, width * sizeof(uchar3), height);
    uchar3* d_img2; size_t pitch2;
    cudaMallocPitch(&d_img2, &pitch2, width * sizeof(uchar3), height);

    MouseParams mouse_state;                                            // kernel.cu:255-257
    mouse_state.pose = &camera.pose;
    float angle = 0.0f;
    double fps = 0.0;
    for (int l = 0; l < frames; l++) {
        angle += 0.005f;
        b.pose.yaw = 0.6f + angle;                                      // the animation the reference sketches (kernel.cu:272-273)
        scene.update_mesh_instance(1, b);
        auto t0 = std::chrono::steady_clock::now();
        camera.render_scene(scene, d_img, pitch);
        camera.render_scene(scene, d_img2, pitch2);
        cudaDeviceSynchronize();
        auto err = cudaGetLastError();
        if (err != cudaSuccess) std::cerr << "CUDA error: " << cudaGetErrorString(err) << std::endl;
        fps = 1.0 / std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (l % 10 == 0 || l == frames - 1) std::cout << "FPS: " << fps << "\n";
        if (present) {
            display_image(d_img, width, height, pitch, fps, mouse_state);  // kernel.cu:288
            std::cout << "pose: "; transforms::print(camera.pose);
        }
    }
    if (!present) write_ppm(out, d_img, width, height, pitch);
    cudaFree(d_img);
    cudaFree(d_img2);
    return 0;
}
