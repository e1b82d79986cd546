// raytracer_scene.hpp -- Scene, Camera and OBJLoader of the drop-in surface (included by raytracer.hpp).
#pragma once

// Scene.h:10-30 / Scene.cpp:4-74.  add_* copy by value; upload_to_device() builds every BLAS on the GPU, uploads
// materials (and their texels) and builds the TLAS; update_mesh_instance() re-uploads one record and re-fits the TLAS.
class Scene {
    std::vector<Material> materials;
    std::vector<MeshPrimitive> meshes;
    std::vector<MeshInstance> mesh_instances;
    std::shared_ptr<rtb_detail::Context> context;

public:
    Scene() : d_materials(nullptr), d_meshes(nullptr), d_mesh_instances(nullptr), num_mesh_instances(0) {}

    void add_material(Material material) { materials.push_back(material); }
    void add_mesh(MeshPrimitive mesh) { meshes.push_back(mesh); }
    void add_mesh_instance(MeshInstance mesh_instance) { mesh_instances.push_back(mesh_instance); }

    Material* d_materials;              // opaque non-null tokens after upload_to_device() (kept for source compatibility)
    d_MeshPrimitive* d_meshes;
    MeshInstance* d_mesh_instances;
    int num_mesh_instances;

    rtb_context* handle()
    {
        if (!context) {
            context = std::make_shared<rtb_detail::Context>();
            int device = 0;
            cudaGetDevice(&device);
            if (rtb_context_create(device, &context->ctx) != RTB_OK) rtb_detail::die("Scene: cannot create a GPU context");
        }
        return context->ctx;
    }

    void upload_to_device()
    {
        rtb_context* ctx = handle();
        for (auto& m : meshes) {
            auto& st = *m.bvh_top.state;
            if (st.mesh_id >= 0 && st.owner == context) continue;          // already resident (shared by value copies)
            const rtb_triangle* t = reinterpret_cast<const rtb_triangle*>(st.triangles.data());
            if (rtb_mesh_create(ctx, t, (int)st.triangles.size(), 0, &st.mesh_id) != RTB_OK) rtb_detail::die("Scene::upload_to_device (mesh)");
            st.owner = context;
        }
        std::vector<rtb_material> mats(materials.size());
        for (size_t i = 0; i < materials.size(); i++) {
            Material& m = materials[i];
            if (m.texels && !m.texels->uploaded) {
                unsigned char* dev = nullptr;
                if (rtb_texture_upload(ctx, m.texels->bgr.data(), m.texels->w, m.texels->h, &dev, &m.texture_pitch) != RTB_OK)
                    rtb_detail::die("Scene::upload_to_device (texture)");
                m.texture = reinterpret_cast<uchar3*>(dev);
                m.texels->uploaded = true;
            }
            std::memcpy(&mats[i], &m, sizeof(rtb_material));                 // first 48 bytes are the reference's Material
        }
        if (rtb_set_materials(ctx, mats.data(), (int)mats.size()) != RTB_OK) rtb_detail::die("Scene::upload_to_device (materials)");
        std::vector<rtb_instance> inst(mesh_instances.size());
        for (size_t i = 0; i < mesh_instances.size(); i++) {
            mesh_instances[i].build_inv();                                   // Scene.cpp:58
            inst[i] = to_record(mesh_instances[i]);
        }
        if (rtb_set_instances(ctx, inst.data(), (int)inst.size()) != RTB_OK) rtb_detail::die("Scene::upload_to_device (instances)");
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
        if (rtb_update_instance(handle(), index, &rec) != RTB_OK) rtb_detail::die("Scene::update_mesh_instance");
    }

    // per-frame BLAS rebuild for deforming geometry (config C5; the reference has no such path)
    void rebuild_mesh(int mesh_index, const std::vector<TrianglePrimitive>& triangles)
    {
        auto& st = *meshes[(size_t)mesh_index].bvh_top.state;
        st.triangles = triangles;
        if (rtb_mesh_rebuild(handle(), st.mesh_id, reinterpret_cast<const rtb_triangle*>(triangles.data()), (int)triangles.size(), 0) != RTB_OK)
            rtb_detail::die("Scene::rebuild_mesh");
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

    // img_ptr: caller-owned DEVICE buffer from cudaMallocPitch (kernel.cu:247-253), BGR.  Returns after enqueue
    // on the default stream unless synchronize is set, exactly like the reference.
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
        p.tile_rank = 0; p.tile_count = 1;
        if (rtb_render(scene.handle(), &p, reinterpret_cast<unsigned char*>(img_ptr), pitch, nullptr, nullptr) != RTB_OK)
            std::cerr << "CUDA error: " << rtb_last_error() << std::endl;   // the reference only prints (kernel.cu:281-286)
        if (synchronize) cudaDeviceSynchronize();
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
