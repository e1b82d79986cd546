// raytracer_scene.hpp -- Scene, Camera and OBJLoader of the drop-in surface (included by raytracer.hpp).
#pragma once

// Scene.h:10-30 / Scene.cpp:4-74.  add_* copy by value; upload_to_device() builds every BLAS on the GPU, uploads
// materials (and their texels) and builds the TLAS; update_mesh_instance() re-uploads one record and rebuilds the (small)
// TLAS on the device.
//
// Multi-GPU (SURVEY 8b "Threading / streams"): a Scene can be replicated on N devices of one node -- use_devices(N), or the
// environment variable RTB_DEVICES=N for an application that is not recompiled -- and Camera::render_scene then drives all of
// them from the calling thread: device d renders the 8-row bands b with b % N == d straight into the caller's frame on device 0
// (peer access over NVLink, 16-byte stores), each on its own stream, and device 0's default stream waits for all of them, so the
// application's existing cudaDeviceSynchronize() still means "the frame is complete".  One rtb_context per device.
class Scene {
    std::vector<Material> materials;
    std::vector<MeshPrimitive> meshes;
    std::vector<MeshInstance> mesh_instances;
    std::vector<std::shared_ptr<rtb_detail::Context>> contexts;      // one per device in use; [0] renders into the caller's frame locally
    int want_devices = 0;

    struct DeviceState { int device = 0; cudaStream_t stream = nullptr; cudaEvent_t done = nullptr; };
    std::vector<DeviceState> devs;

    void open_contexts()
    {
        if (!contexts.empty()) return;
        int first = 0, count = 1;
        cudaGetDevice(&first);
        cudaGetDeviceCount(&count);
        int n = want_devices;
        if (n <= 0) { const char* e = std::getenv("RTB_DEVICES"); n = e ? std::atoi(e) : 1; }
        if (n < 1) n = 1;
        for (int d = 0; d < n; d++) {
            // more contexts than devices wrap around (two contexts on one device: what the single-GPU tests exercise)
            const int device = (first + d) % (count > 0 ? count : 1);
            auto c = std::make_shared<rtb_detail::Context>();
            if (rtb_context_create(device, &c->ctx) != RTB_OK) rtb_detail::die("Scene: cannot create a GPU context");
            contexts.push_back(c);
            DeviceState st;
            st.device = device;
            if (d > 0) {
                cudaSetDevice(device);
                if (device != first) {
                    cudaError_t e = cudaDeviceEnablePeerAccess(first, 0);        // this device's kernels store into device `first`'s frame
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { std::cerr << "Scene: no peer access from device " << device << " to " << first << std::endl; std::exit(1); }
                    cudaGetLastError();
                }
                cudaStreamCreateWithFlags(&st.stream, cudaStreamNonBlocking);
                cudaEventCreateWithFlags(&st.done, cudaEventDisableTiming);
            }
            devs.push_back(st);
        }
        cudaSetDevice(first);
    }

public:
    Scene() : d_materials(nullptr), d_meshes(nullptr), d_mesh_instances(nullptr), num_mesh_instances(0) {}

    void add_material(Material material) { materials.push_back(material); }
    void add_mesh(MeshPrimitive mesh) { meshes.push_back(mesh); }
    void add_mesh_instance(MeshInstance mesh_instance) { mesh_instances.push_back(mesh_instance); }

    Material* d_materials;              // opaque non-null tokens after upload_to_device() (kept for source compatibility)
    d_MeshPrimitive* d_meshes;
    MeshInstance* d_mesh_instances;
    int num_mesh_instances;

    // replicate the scene on the first n devices (call before upload_to_device; default: RTB_DEVICES or 1)
    void use_devices(int n) { if (contexts.empty()) want_devices = n; }
    int num_devices() { open_contexts(); return (int)contexts.size(); }
    rtb_context* handle(int d = 0) { open_contexts(); return contexts[(size_t)d]->ctx; }
    int device_of(int d) { open_contexts(); return devs[(size_t)d].device; }
    cudaStream_t stream_of(int d) { open_contexts(); return devs[(size_t)d].stream; }
    cudaEvent_t event_of(int d) { open_contexts(); return devs[(size_t)d].done; }

    void upload_to_device()
    {
        open_contexts();
        int prev = 0;
        cudaGetDevice(&prev);
        for (size_t d = 0; d < contexts.size(); d++) {
            rtb_context* ctx = contexts[d]->ctx;
            for (auto& m : meshes) {
                auto& st = *m.bvh_top.state;
                if (st.mesh_id >= 0 && st.owner == contexts[0] && st.replicas > (int)d) continue;   // already resident (shared by value copies)
                const rtb_triangle* t = reinterpret_cast<const rtb_triangle*>(st.triangles.data());
                int id = -1;
                if (rtb_mesh_create(ctx, t, (int)st.triangles.size(), 0, &id) != RTB_OK) rtb_detail::die("Scene::upload_to_device (mesh)");
                if (d == 0) { st.mesh_id = id; st.owner = contexts[0]; }
                else if (id != st.mesh_id) { std::cerr << "Scene: mesh ids differ between devices" << std::endl; std::exit(1); }
                st.replicas = (int)d + 1;
            }
            std::vector<rtb_material> mats(materials.size());
            for (size_t i = 0; i < materials.size(); i++) {
                Material& m = materials[i];
                rtb_material rec;
                std::memcpy(&rec, &m, sizeof(rtb_material));                     // first 48 bytes are the reference's Material
                if (m.texels) {
                    if (m.texels->dev.size() <= d) m.texels->dev.resize(d + 1, nullptr);
                    if (!m.texels->dev[d]) {                                     // texels live once per device
                        size_t tp = 0;
                        if (rtb_texture_upload(ctx, m.texels->bgr.data(), m.texels->w, m.texels->h, &m.texels->dev[d], &tp) != RTB_OK)
                            rtb_detail::die("Scene::upload_to_device (texture)");
                        m.texels->pitch = tp;
                    }
                    if (d == 0) { m.texture = reinterpret_cast<uchar3*>(m.texels->dev[0]); m.texture_pitch = m.texels->pitch; m.texels->uploaded = true; }
                    std::memcpy(&rec, &m, sizeof(rtb_material));
                    rec.texture = m.texels->dev[d];
                    rec.texture_pitch = m.texels->pitch;
                }
                mats[i] = rec;
            }
            if (rtb_set_materials(ctx, mats.data(), (int)mats.size()) != RTB_OK) rtb_detail::die("Scene::upload_to_device (materials)");
            std::vector<rtb_instance> inst(mesh_instances.size());
            for (size_t i = 0; i < mesh_instances.size(); i++) {
                mesh_instances[i].build_inv();                                   // Scene.cpp:58
                inst[i] = to_record(mesh_instances[i]);
            }
            if (rtb_set_instances(ctx, inst.data(), (int)inst.size()) != RTB_OK) rtb_detail::die("Scene::upload_to_device (instances)");
        }
        cudaSetDevice(prev);
        num_mesh_instances = (int)mesh_instances.size();
        d_materials = materials.empty() ? nullptr : materials.data();
        d_mesh_instances = mesh_instances.empty() ? nullptr : mesh_instances.data();
        d_meshes = meshes.empty() ? nullptr : meshes[0].to_device();
    }

    void update_mesh_instance(int index, MeshInstance mesh_instance)       // Scene.cpp:67-74
    {
        mesh_instances[(size_t)index] = mesh_instance;
        mesh_instances[(size_t)index].build_inv();
        rtb_instance rec = to_record(mesh_instances[(size_t)index]);
        open_contexts();
        int prev = 0;
        cudaGetDevice(&prev);
        for (auto& c : contexts)
            if (rtb_update_instance(c->ctx, index, &rec) != RTB_OK) rtb_detail::die("Scene::update_mesh_instance");
        cudaSetDevice(prev);
    }

    // per-frame BLAS rebuild for deforming geometry (config C5; the reference has no such path): every replica rebuilds its own copy
    void rebuild_mesh(int mesh_index, const std::vector<TrianglePrimitive>& triangles)
    {
        auto& st = *meshes[(size_t)mesh_index].bvh_top.state;
        st.triangles = triangles;
        open_contexts();
        int prev = 0;
        cudaGetDevice(&prev);
        for (auto& c : contexts)
            if (rtb_mesh_rebuild(c->ctx, st.mesh_id, reinterpret_cast<const rtb_triangle*>(triangles.data()), (int)triangles.size(), 0) != RTB_OK)
                rtb_detail::die("Scene::rebuild_mesh");
        cudaSetDevice(prev);
    }

private:
    rtb_instance to_record(const MeshInstance& mi)
    {
        rtb_instance r;
        std::memcpy(&r, &mi, sizeof r);
        if (mi.mesh_index < 0 || mi.mesh_index >= (int)meshes.size()) { std::cerr << "MeshInstance: mesh_index out of range" << std::endl; std::exit(1); }
        r.mesh_index = meshes[(size_t)mi.mesh_index].bvh_top.state->mesh_id;   // scene-local index -> context mesh id
        return r;
    }
};

// Camera.h:10-30 / Camera.cu:6-41
class Camera {
public:
    Camera(int width, int height, float3x3 K, float4 D) : width(width), height(height), K(K), D(D), mode(RTB_MODE_REFERENCE), max_depth(0), spp(1), seed(42)
    {
        K_inv = invert_intrinsic(K);
        pose = lre();
    }

    lre pose;
    int width;
    int height;
    float3x3 K;
    float3x3 K_inv;
    float4 D;
    int mode, max_depth, spp;           // extensions: RTB_MODE_*; defaults give the reference's image
    unsigned seed;

    // img_ptr: caller-owned DEVICE buffer from cudaMallocPitch (kernel.cu:247-253), BGR, on the device that was current when the
    // Scene was uploaded.  Returns after enqueue unless synchronize is set, exactly like the reference; with a Scene replicated
    // on N devices every device renders its bands into that buffer and the current device's default stream waits for them.
    void render_scene(Scene& scene, uchar3* img_ptr, size_t pitch, bool synchronize = false)
    {
        rtb_render_params p;
        std::memset(&p, 0, sizeof p);
        p.width = width; p.height = height;
        std::memcpy(p.K_inv, &K_inv.m[0][0], sizeof p.K_inv);
        p.D[0] = D.x; p.D[1] = D.y; p.D[2] = D.z; p.D[3] = D.w;
        p.pose = transforms::detail::to_c(pose);
        p.inv_pose = transforms::detail::to_c(invert_lre(pose));             // Camera.cu:21
        p.mode = mode; p.max_depth = max_depth; p.spp = spp; p.seed = seed;
        const int n = scene.num_devices();
        p.tile_count = n;
        int prev = 0;
        cudaGetDevice(&prev);
        for (int d = n - 1; d >= 0; d--) {                                   // the helpers first, the frame's own device last
            p.tile_rank = d;
            if (rtb_render(scene.handle(d), &p, reinterpret_cast<unsigned char*>(img_ptr), pitch, nullptr, (void*)scene.stream_of(d)) != RTB_OK)
                std::cerr << "CUDA error: " << rtb_last_error() << std::endl;   // the reference only prints (kernel.cu:281-286)
            if (d > 0) cudaEventRecord(scene.event_of(d), scene.stream_of(d));
        }
        cudaSetDevice(scene.device_of(0));
        for (int d = 1; d < n; d++) cudaStreamWaitEvent(nullptr, scene.event_of(d), 0);   // "the frame" = the default stream, as before
        if (synchronize) cudaDeviceSynchronize();
        cudaSetDevice(prev);
    }
};

// OBJLoader.hpp:12-180
namespace OBJLoader {
static MeshPrimitive load(std::string fp)
{
    std::cout << "Loading OBJ file: " << fp << std::endl;
    rtb_triangle* tris = nullptr;
    int n = 0;
    if (rtb_obj_load(fp.c_str(), &tris, &n) != RTB_OK) {
        std::cout << "Could not open file " << fp << std::endl;             // OBJLoader.hpp:23-27
        std::exit(1);
    }
    std::vector<TrianglePrimitive> triangles((size_t)n);
    if (n) std::memcpy((void*)triangles.data(), tris, sizeof(rtb_triangle) * (size_t)n);
    rtb_free(tris);
    std::cout << "OBJ File: " << fp << std::endl;
    std::cout << "Loaded " << triangles.size() << " triangles" << std::endl;
    return MeshPrimitive(triangles);
}
}  // namespace OBJLoader
