// wost_run_config: the elaina-exec configuration runner on top of the C ABI (host only, no CUDA in this file).
//
// Drop-in for run_expr (exec.cu:39-221): same JSON keys, same output directory layout
//   <base_path>/<exp_name>/{conf.json, result.json, <file_name>.exr, <file_name>.png}
// Scene keys follow Problem<DIM>::loadConfig (core/problem.cu:151-213): aabb/{min,max}, evaluation_grid/mData/{scale,pos,
// up[,right]}, mesh/{dirichlet_path,neumann_path,vertex_color_dirichlet_path,vertex_color_neumann_path}, mask_path,
// source_path, {source,dirichlet,neumann}_intensity. Differences, all additive:
//   * mask_path: binary PGM/PPM (the reference decodes any stb format); absent -> every pixel, at any frame size
//     (the reference allocates a fixed 1024^2 mask, core/problem.cu:245-247);
//   * source term: "source_dense": {"path": raw float32 file, "dims": [nx,ny,nz], "origin": [x,y,z], "voxel": h} -- a
//     dense Vec3f grid; NanoVDB files (source_path) are SURVEY 8(f) row N3 and return WOST_ERR_UNSUPPORTED;
//   * integrator.setting may carry "seed", "sppBatch"; result.json also reports walk_steps and walk_steps_per_sec;
//   * export type "raw" writes <file_name>.f32, the fp32 film {r,g,b,a,weight} x W*H unflipped (the .exr/.png pair is for viewing);
//   * errors come back as wost_status (reference: exit(1) / uncaught std::runtime_error).
#include <sys/stat.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <fstream>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#include "../../include/wost.h"
#include "imageio.h"
#include "json.h"
#include "nvdb.h"

namespace wost
{
    void set_last_error(const std::string &msg);
}

namespace
{
    struct RunError
    {
        wost_status code;
        std::string msg;
    };
    [[noreturn]] void fail(wost_status c, const std::string &m) { throw RunError{c, m}; }
    void ck(wost_status s)
    {
        if (s != WOST_OK)
            fail(s, wost_last_error());
    }
    bool read_file(const std::string &path, std::string &out)
    {
        std::ifstream f(path, std::ios::binary);
        if (!f.is_open())
            return false;
        std::stringstream ss;
        ss << f.rdbuf();
        out = ss.str();
        return true;
    }
    void mkdirs(const std::string &path)
    {
        std::string cur;
        for (size_t i = 0; i <= path.size(); ++i)
        {
            if (i == path.size() || path[i] == '/')
            {
                if (!cur.empty())
                    mkdir(cur.c_str(), 0777);
            }
            if (i < path.size())
                cur += path[i];
        }
    }
    // json_get_or_throw with '/'-separated nested paths (core/common.h:152-171)
    const wjson::Value &get(const wjson::Value &v, const std::string &path)
    {
        const wjson::Value *cur = &v;
        size_t pos = 0;
        while (pos <= path.size())
        {
            const size_t nx = path.find('/', pos);
            const std::string key = path.substr(pos, nx == std::string::npos ? std::string::npos : nx - pos);
            cur = cur->find(key);
            if (!cur)
                fail(WOST_ERR_CONFIG, "Key not found: " + path);
            if (nx == std::string::npos)
                break;
            pos = nx + 1;
        }
        return *cur;
    }
    const wjson::Value *opt(const wjson::Value &v, const std::string &key) { return v.find(key); }
    void get_vec(const wjson::Value &v, const std::string &path, int n, float *out)
    {
        const wjson::Value &a = get(v, path);
        if (!a.is_array() || (int)a.size() < n)
            fail(WOST_ERR_CONFIG, "expected an array of " + std::to_string(n) + " numbers at " + path);
        for (int i = 0; i < n; ++i)
            out[i] = a.at((size_t)i).as_float();
    }

    // lbvh::scene_loader<DIM> (ext/lbvh/include/snch_lbvh/scene_loader.cuh:49-82, 118-151): "v x y z", "l i j" / "f i j k"
    void load_obj(const std::string &path, int dim, std::vector<float> &verts, std::vector<int> &prims)
    {
        std::ifstream f(path);
        if (!f.is_open())
            fail(WOST_ERR_IO, "Could not open .obj file: " + path);
        std::string line;
        while (std::getline(f, line))
        {
            std::istringstream ls(line);
            std::string prefix;
            ls >> prefix;
            if (prefix == "v")
            {
                float c[3] = {0, 0, 0};
                ls >> c[0] >> c[1] >> c[2];
                for (int k = 0; k < dim; ++k)
                    verts.push_back(c[k]);
            }
            else if ((dim == 2 && prefix == "l") || (dim == 3 && prefix == "f"))
            {
                for (int k = 0; k < dim; ++k)
                {
                    int i = 0;
                    ls >> i;
                    prims.push_back(i - 1);
                }
            }
        }
    }
    // parseVertexColorFile (core/problem.cu:63-96)
    void load_colors(const std::string &path, std::vector<float> &out)
    {
        std::string text;
        if (!read_file(path, text))
            fail(WOST_ERR_IO, "Could not open vertex colour file: " + path);
        const wjson::Value doc = wjson::parse(text);
        const wjson::Value &arr = get(doc, "ColorConfigurations");
        if (!arr.is_array())
            fail(WOST_ERR_CONFIG, "The ColorConfigurations item is not an array.");
        for (size_t i = 0; i < arr.size(); ++i)
        {
            const wjson::Value &c = arr.at(i);
            if (get(c, "vertexID").as_int() != (int)i + 1)
                fail(WOST_ERR_CONFIG, "The configurations should be sorted.");
            const char *keys[6] = {"leftColor/R", "leftColor/G", "leftColor/B", "rightColor/R", "rightColor/G", "rightColor/B"};
            for (const char *k : keys)
                out.push_back(get(c, k).as_float());
        }
    }

    // ---- tone maps of exportEnergy (util/film.h:109-144, util/tonemapping.cuh). The reference evaluates 256-entry
    // LUTs; these are piecewise-linear fits through the maps' published anchor colours.
    void lerp_map(const float (*pts)[3], int n, float t, float rgb[3])
    {
        t = std::isnan(t) ? 0.0f : std::min(1.0f, std::max(0.0f, t));
        const float x = t * (float)(n - 1);
        const int i = std::min(n - 2, (int)x);
        const float f = x - (float)i;
        for (int c = 0; c < 3; ++c)
            rgb[c] = pts[i][c] + f * (pts[i + 1][c] - pts[i][c]);
    }
    void tone(const std::string &name, float v, float vn, float rgb[3])
    {
        static const float jet[9][3] = {{0, 0, 0.5f}, {0, 0, 1}, {0, 0.5f, 1}, {0, 1, 1}, {0.5f, 1, 0.5f}, {1, 1, 0}, {1, 0.5f, 0}, {1, 0, 0}, {0.5f, 0, 0}};
        static const float parula[9][3] = {{0.2422f, 0.1504f, 0.6603f}, {0.2810f, 0.3228f, 0.9579f}, {0.1786f, 0.5289f, 0.9682f},
                                           {0.0689f, 0.6948f, 0.8394f}, {0.2161f, 0.7843f, 0.5923f}, {0.6720f, 0.7793f, 0.2227f},
                                           {0.9970f, 0.7659f, 0.2199f}, {0.9769f, 0.9029f, 0.1573f}, {0.9769f, 0.9839f, 0.0805f}};
        static const float rdbu[11][3] = {{0.4039f, 0, 0.1216f}, {0.6980f, 0.0941f, 0.1686f}, {0.8392f, 0.3765f, 0.3020f}, {0.9569f, 0.6471f, 0.5098f},
                                          {0.9922f, 0.8588f, 0.7804f}, {0.9686f, 0.9686f, 0.9686f}, {0.8196f, 0.8980f, 0.9412f},
                                          {0.5725f, 0.7725f, 0.8706f}, {0.2627f, 0.5765f, 0.7647f}, {0.1294f, 0.4000f, 0.6745f}, {0.0196f, 0.1882f, 0.3804f}};
        if (name == "NONE_NORMALIZED")
            rgb[0] = rgb[1] = rgb[2] = vn;
        else if (name == "MATLAB_JET")
            lerp_map(jet, 9, vn, rgb);
        else if (name == "MATLAB_PARULA")
            lerp_map(parula, 9, vn, rgb);
        else if (name == "IDL_RDBU")
            lerp_map(rdbu, 11, vn, rgb);
        else
            rgb[0] = rgb[1] = rgb[2] = v;
    }
    // Film::save / Film::saveEnergy + Image::saveImage: <stem>.exr and <stem>.png, vertically flipped
    void save_pair(const std::string &stem, int w, int h, const std::vector<float> &rgba)
    {
        if (!wimg::write_exr(stem + ".exr", w, h, rgba.data(), true))
            fail(WOST_ERR_IO, "cannot write " + stem + ".exr");
        std::vector<uint8_t> px(rgba.size());
        for (size_t i = 0; i < rgba.size(); ++i)
        {
            const float v = rgba[i] * 255.0f;
            const int q = std::isnan(v) ? 0 : (v <= 0.0f ? 0 : (v >= 255.0f ? 255 : (int)v));
            px[i] = (uint8_t)q;
        }
        if (!wimg::write_png(stem + ".png", w, h, px.data(), true))
            fail(WOST_ERR_IO, "cannot write " + stem + ".png");
    }
    int channel_of(const std::string &name)
    {
        if (name == "SOLUTION")
            return WOST_CH_SOLUTION;
        if (name == "DIRICHLET_SDF")
            return WOST_CH_DIRICHLET_SDF;
        if (name == "NEUMANN_SDF")
            return WOST_CH_NEUMANN_SDF;
        if (name == "SOURCE")
            return WOST_CH_SOURCE;
        return -1;
    }

    struct Handles
    {
        wost_scene *dir = nullptr, *neu = nullptr;
        wost_problem *prob = nullptr;
        wost_integrator *it = nullptr;
        wost_comm *comm = nullptr; // non-null when started by torchrun / mpirun with more than one process (one per GPU)
        ~Handles()
        {
            wost_integrator_destroy(it);
            wost_problem_destroy(prob);
            wost_scene_destroy(dir);
            wost_scene_destroy(neu);
            wost_comm_destroy(comm);
        }
    };

    void run(const std::string &conf_path)
    {
        std::string text;
        if (!read_file(conf_path, text))
            fail(WOST_ERR_IO, "Configuration file does not exist: " + conf_path);
        const wjson::Value conf = wjson::parse(text);
        const int dim = get(conf, "dimensionality").as_int();
        const std::string base = get(conf, "base_path").as_string(), exp = get(conf, "exp_name").as_string();
        const std::string out_dir = (base.empty() || base.back() == '/' ? base : base + "/") + exp;
        // Multi-GPU (no counterpart in the reference, which is single-device): under torchrun / mpirun every process takes the
        // GPU LOCAL_RANK, evaluates its interleaved row tiles of the frame, joins the gradient all-reduce of the guiding network,
        // and rank 0 gathers the film and writes the outputs. A single process behaves exactly like the reference.
        Handles H;
        ck(wost_comm_create_from_env(&H.comm));
        int rank = 0, world = 1;
        wost_comm_info(H.comm, &rank, &world, nullptr);
        const bool writer = rank == 0;
        mkdirs(out_dir);
        if (writer)
        {
            std::string dumped;
            wjson::dump(conf, dumped, 4);
            std::ofstream f(out_dir + "/conf.json");
            f << dumped << std::endl;
        }
        const wjson::Value &scene = get(conf, "scene");
        const wjson::Value &integ = get(conf, "integrator");
        const std::string type = get(integ, "type").as_string();
        const wjson::Value &setting = get(integ, "setting");
        if (dim != 2 && dim != 3)
            fail(WOST_ERR_UNSUPPORTED, "Unsupported dimensionality.");
        if (type != "uniform" && type != "guided")
            fail(WOST_ERR_UNSUPPORTED, "Unrecognized integrator type.");

        // ---- Problem<DIM>::loadConfig
        wost_problem_desc d;
        memset(&d, 0, sizeof(d));
        d.dim = dim;
        get_vec(scene, "aabb/min", dim, d.aabb_min);
        get_vec(scene, "aabb/max", dim, d.aabb_max);
        const wjson::Value &probe = get(scene, "evaluation_grid/mData");
        d.grid_scale = get(probe, "scale").as_float();
        get_vec(probe, "pos", dim, d.grid_pos);
        get_vec(probe, "up", dim, d.grid_up);
        d.grid_right[0] = 1.0f;
        if (dim == 3)
            get_vec(probe, "right", 3, d.grid_right);
        const wjson::Value &mesh = get(scene, "mesh");
        std::vector<float> dverts, nverts, dcol, ncol, source;
        std::vector<int> dprims, nprims;
        if (const wjson::Value *p = opt(mesh, "dirichlet_path"))
        {
            load_obj(p->as_string(), dim, dverts, dprims);
            ck(wost_scene_create(dim, dverts.data(), (int)dverts.size() / dim, dprims.data(), (int)dprims.size() / dim, &H.dir));
        }
        if (const wjson::Value *p = opt(mesh, "neumann_path"))
        {
            load_obj(p->as_string(), dim, nverts, nprims);
            ck(wost_scene_create(dim, nverts.data(), (int)nverts.size() / dim, nprims.data(), (int)nprims.size() / dim, &H.neu));
        }
        d.dirichlet = H.dir;
        d.neumann = H.neu;
        const wjson::Value *pc;
        if (H.dir && (pc = opt(mesh, "vertex_color_dirichlet_path")))
        {
            load_colors(pc->as_string(), dcol);
            d.color_dirichlet = dcol.data();
            d.n_color_dirichlet = (int)dcol.size() / 6;
        }
        if (H.neu && (pc = opt(mesh, "vertex_color_neumann_path")))
        {
            load_colors(pc->as_string(), ncol);
            d.color_neumann = ncol.data();
            d.n_color_neumann = (int)ncol.size() / 6;
        }
        d.source_intensity = (float)scene.number_or("source_intensity", 1.0);
        d.dirichlet_intensity = (float)scene.number_or("dirichlet_intensity", 1.0);
        d.neumann_intensity = (float)scene.number_or("neumann_intensity", 1.0);
        if (const wjson::Value *sp = opt(scene, "source_path"))
        {
            // core/problem.cu:135-149: a NanoVDB Vec3f grid, flattened over its index bounding box (nvdb.h)
            wnvdb::DenseVec3f g;
            const std::string err = wnvdb::read_vec3f(sp->as_string(), g);
            if (!err.empty())
                fail(err.rfind("cannot", 0) == 0 ? WOST_ERR_IO : WOST_ERR_UNSUPPORTED, err);
            for (int k = 0; k < 3; ++k)
            {
                d.source_dims[k] = g.dims[k];
                d.source_origin[k] = g.origin[k];
            }
            d.source_voxel = g.voxel;
            source.swap(g.data);
            d.source = source.data();
        }
        else if (const wjson::Value *sd = opt(scene, "source_dense"))
        {
            float dims[3];
            get_vec(*sd, "dims", 3, dims);
            get_vec(*sd, "origin", 3, d.source_origin);
            d.source_voxel = get(*sd, "voxel").as_float();
            for (int k = 0; k < 3; ++k)
                d.source_dims[k] = (int)dims[k];
            std::string raw;
            if (!read_file(get(*sd, "path").as_string(), raw))
                fail(WOST_ERR_IO, "cannot read the dense source grid");
            const size_t n = (size_t)d.source_dims[0] * d.source_dims[1] * d.source_dims[2] * 3;
            if (raw.size() != n * sizeof(float))
                fail(WOST_ERR_CONFIG, "source_dense: file size does not match dims");
            source.resize(n);
            memcpy(source.data(), raw.data(), raw.size());
            d.source = source.data();
        }
        std::vector<uint8_t> mask;
        if (const wjson::Value *mp = opt(scene, "mask_path"))
        {
            if (!wimg::read_pnm_mask(mp->as_string(), d.mask_w, d.mask_h, mask))
                fail(WOST_ERR_IO, "mask_path: only binary PGM/PPM masks are read by this build");
            // the reference loads the mask flipped (core/problem.cu:222: loadImage(path, flip = true))
            std::vector<uint8_t> fl(mask.size());
            for (int y = 0; y < d.mask_h; ++y)
                memcpy(&fl[(size_t)y * d.mask_w], &mask[(size_t)(d.mask_h - 1 - y) * d.mask_w], (size_t)d.mask_w);
            mask.swap(fl);
            d.mask = mask.data();
        }
        ck(wost_problem_create(&d, &H.prob));

        // ---- integrator (exec.cu:86-123)
        wost_settings s;
        memset(&s, 0, sizeof(s));
        float fs[2];
        get_vec(setting, "frameSize", 2, fs);
        s.frame_w = (int)fs[0], s.frame_h = (int)fs[1];
        s.samples_per_pixel = get(setting, "samplesPerPixel").as_int();
        s.max_walking_depth = get(setting, "maxWalkingDepth").as_int();
        s.epsilon_shell = get(setting, "epsilonShell").as_float();
        s.debug_pixel = (int)setting.number_or("debugPixel", 0);
        s.save_spp_metrics_duration = (int)setting.number_or("saveSppMetricsDuration", -1);
        s.save_spp_metrics_until = (int)setting.number_or("saveSppMetricsUntil", 1024);
        s.save_time_metrics_duration = (int)setting.number_or("saveTimeMetricsDuration", -1);
        s.seed = (uint32_t)setting.number_or("seed", 0);
        s.spp_batch = (int)setting.number_or("sppBatch", 0);
        const bool guided = type == "guided";
        if (guided)
        {
            s.train_spp_count = get(setting, "trainSppCount").as_int();
            s.uniform_fraction_training = get(setting, "uniformFractionInTrainingPhase").as_float();
            s.uniform_fraction_guiding = get(setting, "uniformFractionInGuidingPhase").as_float();
            s.max_guided_depth_training = get(setting, "maxGuidedDepthInTrainingPhase").as_int();
            s.max_guided_depth_guiding = get(setting, "maxGuidedDepthInGuidingPhase").as_int();
        }
        ck(wost_comm_shard_settings(H.comm, &s, 4));
        ck(wost_integrator_create(H.prob, &s, guided ? WOST_GUIDED : WOST_UNIFORM, &H.it));
        if (guided)
        {
            std::string net;
            wjson::dump(get(conf, "network"), net, 0);
            ck(wost_reset_network(H.it, net.c_str()));
            ck(wost_integrator_set_comm(H.it, H.comm));
        }

        // ---- channels (exec.cu:131-174)
        std::set<int> channels;
        const wjson::Value &chs = get(integ, "channels");
        for (size_t i = 0; i < chs.size(); ++i)
        {
            const int c = channel_of(chs.at(i).as_string());
            if (c < 0)
                fail(WOST_ERR_CONFIG, "Unrecognized integrator channel: " + chs.at(i).as_string());
            channels.insert(c);
        }
        // progressive metrics (uniform/integrator.cu:577-609): frames/<sampleId>.{exr,png}, frames_time/<ms>.{exr,png}
        struct FrameSink
        {
            std::string dir;
        } sink{out_dir};
        if (world == 1 && (s.save_spp_metrics_duration > 0 || s.save_time_metrics_duration > 0)) // progressive dumps: single process only
        {
            if (s.save_spp_metrics_duration > 0)
                mkdirs(out_dir + "/frames");
            if (s.save_time_metrics_duration > 0)
                mkdirs(out_dir + "/frames_time");
            ck(wost_set_frame_callback(
                H.it,
                [](void *user, int kind, int sample_id, uint64_t elapsed_ms, const float *film, int w, int h)
                {
                    const FrameSink *fs = static_cast<const FrameSink *>(user);
                    std::vector<float> rgba((size_t)w * h * 4);
                    for (size_t p = 0; p < (size_t)w * h; ++p)
                        for (int k = 0; k < 4; ++k)
                            rgba[p * 4 + k] = film[p * 5 + k] / film[p * 5 + 4]; // Film::save: pixel / weight
                    const std::string stem = kind == 0 ? fs->dir + "/frames/" + std::to_string(sample_id) : fs->dir + "/frames_time/" + std::to_string(elapsed_ms);
                    try
                    {
                        save_pair(stem, w, h, rgba);
                    }
                    catch (...)
                    {
                    }
                },
                &sink));
        }
        uint64_t duration = 0, steps = 0;
        float dev_ms = 0.0f;
        bool solved = false;
        for (int c : channels)
        {
            if (c == WOST_CH_SOLUTION)
            {
                ck(wost_solve(H.it, &duration, &steps, &dev_ms));
                ck(wost_film_gather(H.it, H.comm, WOST_CH_SOLUTION, 0));
                double tmax[2] = {(double)duration, (double)dev_ms}, ssum[1] = {(double)steps};
                ck(wost_comm_allreduce_f64(H.comm, tmax, 2, 1)); // whole-job numbers: time = max over ranks, walk-steps = sum
                ck(wost_comm_allreduce_f64(H.comm, ssum, 1, 0));
                duration = (uint64_t)tmax[0], dev_ms = (float)tmax[1], steps = (uint64_t)ssum[0];
                solved = true;
            }
            else
                ck(wost_render_channel(H.it, c));
        }
        if (!writer)
            return;
        if (guided && conf.number_or("print_network", 0) != 0)
        {
            float p[3] = {0.0f, dim == 3 ? -0.21f : 0.0f, 0.0f}, out[48];
            ck(wost_query_network(H.it, p, out));
            for (int i = 0; i < 8; ++i)
                printf("Component %d: lambda_raw = %f, kappa_raw = %f, mu = (%f, %f)\n", i, out[i * (dim + 2)], out[i * (dim + 2) + 1], out[i * (dim + 2) + 2],
                       out[i * (dim + 2) + 3]);
        }

        // ---- export (exec.cu:187-214)
        const wjson::Value &exports = get(conf, "export");
        std::vector<float> film((size_t)s.frame_w * s.frame_h * 5), rgba((size_t)s.frame_w * s.frame_h * 4);
        for (size_t i = 0; i < exports.size(); ++i)
        {
            const wjson::Value &e = exports.at(i);
            const std::string etype = get(e, "type").as_string(), file_name = get(e, "file_name").as_string();
            const int c = channel_of(get(e, "channel").as_string());
            if (c < 0)
                fail(WOST_ERR_CONFIG, "Unrecognized export channel: " + get(e, "channel").as_string());
            ck(wost_film_download(H.it, c, film.data()));
            const size_t np = (size_t)s.frame_w * s.frame_h;
            if (etype == "image")
            {
                for (size_t p = 0; p < np; ++p)
                    for (int k = 0; k < 4; ++k)
                        rgba[p * 4 + k] = film[p * 5 + k] / film[p * 5 + 4]; // Film::save: pixel / weight
            }
            else if (etype == "energy")
            {
                const std::string tname = get(e, "tone").as_string();
                if (tname != "NONE" && tname != "NONE_NORMALIZED" && tname != "MATLAB_JET" && tname != "MATLAB_PARULA" && tname != "IDL_RDBU")
                    continue; // reference: "Unrecognized tone mapping method, skipping..."
                float lo = INFINITY, hi = -INFINITY;
                std::vector<float> lum(np);
                for (size_t p = 0; p < np; ++p)
                {
                    lum[p] = film[p * 5] * 0.299f + film[p * 5 + 1] * 0.587f + film[p * 5 + 2] * 0.114f; // saveEnergy: not divided by weight
                    lo = std::min(lo, lum[p]);
                    hi = std::max(hi, lum[p]);
                }
                for (size_t p = 0; p < np; ++p)
                {
                    tone(tname, lum[p], (lum[p] - lo) / (hi - lo), &rgba[p * 4]);
                    rgba[p * 4 + 3] = 1.0f;
                }
            }
            else if (etype == "raw")
            {
                // extension (the reference skips export types it does not know): the fp32 film {r,g,b,a,weight} x W*H, unflipped
                std::ofstream rf(out_dir + "/" + file_name + ".f32", std::ios::binary);
                rf.write(reinterpret_cast<const char *>(film.data()), (std::streamsize)(film.size() * sizeof(float)));
                if (!rf)
                    fail(WOST_ERR_IO, "cannot write " + out_dir + "/" + file_name + ".f32");
                continue;
            }
            else
                continue;
            save_pair(out_dir + "/" + file_name, s.frame_w, s.frame_h, rgba);
        }

        // ---- result.json (exec.cu:217-220)
        wjson::Value res;
        res.type = wjson::Value::Object;
        if (solved)
        {
            res.obj.emplace_back("duration", wjson::make_number((double)duration));
            res.obj.emplace_back("walk_steps", wjson::make_number((double)steps));
            res.obj.emplace_back("walk_steps_per_sec", wjson::make_number(dev_ms > 0 ? (double)steps / (dev_ms * 1e-3) : 0.0));
            res.obj.emplace_back("n_gpus", wjson::make_number((double)world));
        }
        char ts[64];
        const time_t now = time(nullptr);
        strftime(ts, sizeof(ts), "%Y-%m-%d %H:%M:%S", localtime(&now));
        res.obj.emplace_back("timestamp", wjson::make_string(ts));
        std::string dumped;
        wjson::dump(res, dumped, 4);
        std::ofstream f(out_dir + "/result.json");
        f << dumped << std::endl;
    }
} // namespace

extern "C" wost_status wost_nvdb_read_vec3f(const char *path, int dims[3], float origin[3], float *voxel, float *dense, uint64_t capacity_floats)
{
    if (!path || !dims || !origin || !voxel)
    {
        wost::set_last_error("wost_nvdb_read_vec3f: null argument");
        return WOST_ERR_INVALID;
    }
    wnvdb::DenseVec3f g;
    const std::string err = wnvdb::read_vec3f(path, g);
    if (!err.empty())
    {
        wost::set_last_error(err);
        return err.rfind("cannot", 0) == 0 ? WOST_ERR_IO : WOST_ERR_UNSUPPORTED;
    }
    for (int k = 0; k < 3; ++k)
    {
        dims[k] = g.dims[k];
        origin[k] = g.origin[k];
    }
    *voxel = g.voxel;
    if (dense)
    {
        if (capacity_floats < g.data.size())
        {
            wost::set_last_error("wost_nvdb_read_vec3f: buffer too small");
            return WOST_ERR_INVALID;
        }
        memcpy(dense, g.data.data(), g.data.size() * sizeof(float));
    }
    return WOST_OK;
}

extern "C" wost_status wost_run_config(const char *json_path)
{
    if (!json_path)
    {
        wost::set_last_error("wost_run_config: null path");
        return WOST_ERR_INVALID;
    }
    try
    {
        run(json_path);
        return WOST_OK;
    }
    catch (const RunError &e)
    {
        wost::set_last_error(e.msg);
        return e.code;
    }
    catch (const wjson::Error &e)
    {
        wost::set_last_error(e.what());
        return WOST_ERR_CONFIG;
    }
    catch (const std::exception &e)
    {
        wost::set_last_error(e.what());
        return WOST_ERR_INVALID;
    }
}
