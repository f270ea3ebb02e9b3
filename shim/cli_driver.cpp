// A headless command-line driver written against the reference's OWN boundary API only:
//   NetPimpl::TrainingNet / NetPimpl::RuntimeNet   dlib-dnn-pimpl-wrapper/NetPimpl.h:29-102
//   annonet_infer()                                annonet_infer.cpp:32-240 (unmodified source)
//   set_weights()                                  annonet_train.h:20-83
//   tiling::get_tiles()                            tiling/tiling.cpp:23-87
// It exists because the stock annonet_infer_main.cpp cannot be built on a headless box (it pulls in X11 through
// <dlib/gui_widgets.h>) and because parity needs raw tensors, not PNGs.  The same file is compiled twice:
//   shim/Makefile    -> shim/_build/annonet_cli : linked with shim/NetPimpl_b200.cpp + libannonet_b200.so  (the engine)
//   oracle/Makefile  -> oracle/_ref/annonet_ref : linked with the reference's NetPimpl.cpp + dlib's CPU dnn code (the oracle),
//                       with -DAB_WITH_REF_DIRECT adding the verbs that reach behind the pimpl (oracle/ref_direct.cpp)
// All tensors travel as raw little-endian files; shapes are given on the command line.
#include "annonet_infer.h"
#include "annonet_train.h"
#include "tiling/dlib-wrapper.h"
#include "cli_io.h"

#include <chrono>
#include <cstring>
#include <iostream>
#include <sstream>
#include <thread>

#ifdef _OPENMP
#include <omp.h>
#endif

#ifdef AB_WITH_REF_DIRECT
int ref_materialize(const std::string& in_path, const std::string& out_path, unsigned seed);
int ref_grads(const std::string& net_path, const std::vector<NetPimpl::input_type>& images,
              const std::vector<NetPimpl::training_label_type>& labels, const std::string& out_prefix, bool dump_layers);
int ref_forward_layers(const std::string& anet_path, const NetPimpl::input_type& image, const std::string& out_path);
#endif

namespace {

NetPimpl::input_type image_from(const unsigned char* p, long h, long w)
{
    NetPimpl::input_type img(h, w);
    for (long r = 0; r < h; ++r)
        for (long c = 0; c < w; ++c, p += 3)
            img(r, c) = dlib::rgb_pixel(p[0], p[1], p[2]);
    return img;
}

std::vector<NetPimpl::input_type> load_images(const std::string& path, long n, long h, long w)
{
    const auto bytes = ref_io::slurp(path);
    if (static_cast<long>(bytes.size()) != n * h * w * 3) throw std::runtime_error("image file size mismatch: " + path);
    std::vector<NetPimpl::input_type> v;
    for (long i = 0; i < n; ++i)
        v.push_back(image_from(reinterpret_cast<const unsigned char*>(bytes.data()) + i * h * w * 3, h, w));
    return v;
}

std::vector<NetPimpl::training_label_type> load_labels(const std::string& lpath, const std::string& wpath, long n, long h, long w)
{
    const auto lb = ref_io::slurp(lpath);
    const auto wb = ref_io::slurp(wpath);
    if (static_cast<long>(lb.size()) != n * h * w * 2) throw std::runtime_error("label file size mismatch");
    if (static_cast<long>(wb.size()) != n * h * w * 4) throw std::runtime_error("weight file size mismatch");
    const uint16_t* l = reinterpret_cast<const uint16_t*>(lb.data());
    const float* wt = reinterpret_cast<const float*>(wb.data());
    std::vector<NetPimpl::training_label_type> v(n);
    for (long i = 0; i < n; ++i) {
        v[i].set_size(h, w);
        for (long r = 0; r < h; ++r)
            for (long c = 0; c < w; ++c, ++l, ++wt)
                v[i](r, c) = dlib::loss_multiclass_log_per_pixel_weighted_::weighted_label(*l, *wt);
    }
    return v;
}

std::vector<double> parse_doubles(const std::string& s)
{
    std::vector<double> v;
    std::stringstream ss(s);
    std::string item;
    while (std::getline(ss, item, ',')) if (!item.empty()) v.push_back(std::stod(item));
    return v;
}

void write_u16(const std::string& path, const dlib::matrix<uint16_t>& m)
{
    std::ofstream out(path, std::ios::binary);
    for (long r = 0; r < m.nr(); ++r)
        out.write(reinterpret_cast<const char*>(&m(r, 0)), sizeof(uint16_t) * m.nc());
}

int host_threads()
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

double seconds_since(std::chrono::steady_clock::time_point t0)
{
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

void synthetic_image(NetPimpl::input_type& img, long h, long w, unsigned seed)
{
    img.set_size(h, w);
    dlib::rand rnd(seed);
    for (long r = 0; r < h; ++r)
        for (long c = 0; c < w; ++c)
            img(r, c) = dlib::rgb_pixel(rnd.get_random_8bit_number(), rnd.get_random_8bit_number(), rnd.get_random_8bit_number());
}

int usage()
{
    std::cerr << "usage: annonet_ref <config|materialize|runtime|forward|forward_layers|process|infer|grads|train|adam|set_weights|tiles|recommended|bench_infer|bench_train> ...\n";
    return 2;
}

}  // namespace

int main(int argc, char** argv) try
{
    if (argc < 2) return usage();
    const std::string cmd = argv[1];
    auto arg = [&](int i) -> std::string { if (i >= argc) throw std::runtime_error("missing argument"); return argv[i]; };

    if (cmd == "config") {
        // <K> <scaler> <min_filters> <out>   -- same call order as annonet_train_main.cpp:401-405
        NetPimpl::TrainingNet t;
        t.Initialize();
        t.SetNetWidth(std::stod(arg(3)), std::stoi(arg(4)));
        t.SetClassCount(static_cast<unsigned short>(std::stoi(arg(2))));
        t.Serialize(arg(5));
        return 0;
    }
#ifdef AB_WITH_REF_DIRECT
    if (cmd == "materialize") return ref_materialize(arg(2), arg(3), static_cast<unsigned>(std::stoul(arg(4))));
#endif
    if (cmd == "runtime") {
        // <net> <out_anet>   -- RuntimeNet::operator=(TrainingNet), NetPimpl.cpp:226-232
        NetPimpl::TrainingNet t;
        t.Initialize();
        t.Deserialize(arg(2));
        t.GetRuntimeNet().Serialize(arg(3));
        return 0;
    }
    if (cmd == "forward" || cmd == "forward_layers") {
        // <anet> <img> <H> <W> <out>
        const long h = std::stol(arg(4)), w = std::stol(arg(5));
        const auto imgs = load_images(arg(3), 1, h, w);
#ifdef AB_WITH_REF_DIRECT
        if (cmd == "forward_layers") return ref_forward_layers(arg(2), imgs[0], arg(6));
#endif
        NetPimpl::RuntimeNet net;
        net.Deserialize(arg(2));
        const dlib::tensor& out = net.Forward(imgs[0]);
        std::ofstream f(arg(6), std::ios::binary);
        ref_io::write_tensor(f, 1, out);
        return 0;
    }
    if (cmd == "process") {
        // <anet> <img> <H> <W> <out_labels> [gains]
        const long h = std::stol(arg(4)), w = std::stol(arg(5));
        const auto imgs = load_images(arg(3), 1, h, w);
        NetPimpl::RuntimeNet net;
        net.Deserialize(arg(2));
        const std::vector<double> gains = argc > 7 ? parse_doubles(arg(7)) : std::vector<double>();
        const dlib::matrix<uint16_t> labels = net(imgs[0], gains);
        write_u16(arg(6), labels);
        return 0;
    }
    if (cmd == "infer") {
        // <anet> <img> <H> <W> <tile_w> <tile_h> <out_labels> [gains|-] [detection_levels|-] [out_blended_logits]
        const long h = std::stol(arg(4)), w = std::stol(arg(5));
        const auto imgs = load_images(arg(3), 1, h, w);
        NetPimpl::RuntimeNet net;
        net.Deserialize(arg(2));
        tiling::parameters tp;                              // annonet_infer_main.cpp:423-427
        tp.max_tile_width = std::stoi(arg(6));
        tp.max_tile_height = std::stoi(arg(7));
        tp.overlap_x = tp.overlap_y = NetPimpl::TrainingNet::GetRequiredInputDimension();
        const std::vector<double> gains = (argc > 9 && arg(9) != "-") ? parse_doubles(arg(9)) : std::vector<double>();
        const std::vector<double> det = (argc > 10 && arg(10) != "-") ? parse_doubles(arg(10)) : std::vector<double>();
        annonet_infer_temp temp;
        dlib::matrix<uint16_t> result;
        const auto t0 = std::chrono::steady_clock::now();
        annonet_infer(net, imgs[0], result, temp, gains, det, tp);
        const double secs = seconds_since(t0);
        write_u16(arg(8), result);
        if (argc > 11) {
            std::ofstream f(arg(11), std::ios::binary);
            for (const auto& plane : temp.blended_output)
                for (long r = 0; r < plane.nr(); ++r)
                    f.write(reinterpret_cast<const char*>(&plane(r, 0)), sizeof(float) * plane.nc());
        }
        std::cout << "{\"seconds\": " << secs << ", \"mpix_per_s\": " << (h * w / secs / 1e6) << ", \"threads\": " << host_threads() << "}" << std::endl;
        return 0;
    }
    if (cmd == "grads") {
        // <net> <imgs> <labels.u16> <weights.f32> <N> <H> <W> <out_prefix> [layers]
        const long n = std::stol(arg(6)), h = std::stol(arg(7)), w = std::stol(arg(8));
#ifdef AB_WITH_REF_DIRECT
        return ref_grads(arg(2), load_images(arg(3), n, h, w), load_labels(arg(4), arg(5), n, h, w), arg(9), argc > 10);
#else
        throw std::runtime_error("grads needs the reference's internals (oracle build only)");
#endif
    }
    if (cmd == "train") {
        // <net> <imgs> <labels.u16> <weights.f32> <N> <H> <W> <lr> <steps> <out_net> [bn_window] [out_anet] [sync_file]
        // sync_file: TrainingNet::SetSynchronizationFile(name, 0 s) -- dnn_trainer writes its v13 state file (trainer.h:931-964)
        // The files hold steps*N samples; step s consumes samples [s*N, (s+1)*N).
        const long n = std::stol(arg(6)), h = std::stol(arg(7)), w = std::stol(arg(8));
        const double lr = std::stod(arg(9));
        const long steps = std::stol(arg(10));
        const auto imgs = load_images(arg(3), n * steps, h, w);
        const auto labels = load_labels(arg(4), arg(5), n * steps, h, w);
        NetPimpl::TrainingNet t;
        t.Initialize();
        t.Deserialize(arg(2));
        t.SetLearningRate(lr);
        t.SetLearningRateShrinkFactor(1.0);  // no plateau logic inside a parity run
        if (argc > 12 && std::string(arg(12)) != "-")
            t.SetAllBatchNormalizationRunningStatsWindowSizes(std::stoul(arg(12)));
        if (argc > 14) t.SetSynchronizationFile(arg(14), std::chrono::seconds(0));
        const auto t0 = std::chrono::steady_clock::now();
        for (long s = 0; s < steps; ++s) {
            std::vector<NetPimpl::input_type> bi(imgs.begin() + s * n, imgs.begin() + (s + 1) * n);
            std::vector<NetPimpl::training_label_type> bl(labels.begin() + s * n, labels.begin() + (s + 1) * n);
            t.StartTraining(bi, bl);
        }
        NetPimpl::RuntimeNet rt = t.GetRuntimeNet();  // blocks until the trainer thread is idle; cleans the net
        const double secs = seconds_since(t0);
        t.Serialize(arg(11));
        if (argc > 13) rt.Serialize(arg(13));
        std::cout << "{\"seconds\": " << secs << ", \"crops_per_s\": " << (n * steps / secs) << ", \"threads\": " << host_threads()
                  << ", \"learning_rate\": " << t.GetLearningRate() << "}" << std::endl;
        return 0;
    }
#ifdef AB_WITH_REF_DIRECT
    if (cmd == "adam") {
        // <n> <steps> <lr> <wd> <beta1> <beta2> <params.f32> <grads.f32 (steps*n)> <out_params.f32>
        // dlib's own Adam arithmetic (tt::compute_adam_update -> cpu::compute_adam_update, cpu_dlib.cpp:704-743) applied the
        // way adam::operator() + add_layer::update_parameters do (solvers.h:225-260, core.h:1011-1023): t starts at 1.
        const long n = std::stol(arg(2)), steps = std::stol(arg(3));
        const float lr = std::stof(arg(4)), wd = std::stof(arg(5)), b1 = std::stof(arg(6)), b2 = std::stof(arg(7));
        const auto pb = ref_io::slurp(arg(8)), gb = ref_io::slurp(arg(9));
        if (static_cast<long>(pb.size()) != n * 4 || static_cast<long>(gb.size()) != n * steps * 4) throw std::runtime_error("adam: file size mismatch");
        dlib::resizable_tensor params(n), grad(n), m(n), v(n), upd(n);
        std::memcpy(params.host(), pb.data(), pb.size());
        m = 0; v = 0;
        for (long s = 0; s < steps; ++s) {
            std::memcpy(grad.host(), gb.data() + s * n * 4, n * 4);
            dlib::tt::compute_adam_update(0, n, upd, m, v, static_cast<float>(s + 1), lr, wd, b1, b2, params, grad);
            dlib::tt::add(1, params, 1, upd);   // params += step (core.h:1019)
        }
        std::ofstream out(arg(10), std::ios::binary);
        out.write(reinterpret_cast<const char*>(params.host()), n * 4);
        return 0;
    }
#endif
    if (cmd == "set_weights") {
        // <labels.u16> <H> <W> <class_weight> <image_weight> <out_weights.f32>
        const long h = std::stol(arg(3)), w = std::stol(arg(4));
        const auto lb = ref_io::slurp(arg(2));
        if (static_cast<long>(lb.size()) != h * w * 2) throw std::runtime_error("label file size mismatch");
        dlib::matrix<uint16_t> labels(h, w);
        std::memcpy(&labels(0, 0), lb.data(), lb.size());
        NetPimpl::training_label_type weighted;
        set_weights(labels, weighted, std::stod(arg(5)), std::stod(arg(6)));
        std::ofstream out(arg(7), std::ios::binary);
        for (long r = 0; r < h; ++r)
            for (long c = 0; c < w; ++c) {
                const float wt = weighted(r, c).weight;
                out.write(reinterpret_cast<const char*>(&wt), sizeof(wt));
            }
        return 0;
    }
    if (cmd == "tiles") {
        // <W> <H> <max_w> <max_h> <ov_x> <ov_y>: one line per tile, 12 ints (full, non-overlapping, unique as l t r b inclusive)
        tiling::parameters tp;
        tp.max_tile_width = std::stoi(arg(4));
        tp.max_tile_height = std::stoi(arg(5));
        tp.overlap_x = std::stoi(arg(6));
        tp.overlap_y = std::stoi(arg(7));
        for (const auto& t : tiling::get_tiles(std::stoi(arg(2)), std::stoi(arg(3)), tp)) {
            for (const dlib::rectangle* r : {&t.full_rect, &t.non_overlapping_rect, &t.unique_rect})
                std::cout << r->left() << ' ' << r->top() << ' ' << r->right() << ' ' << r->bottom() << ' ';
            std::cout << '\n';
        }
        return 0;
    }
    if (cmd == "recommended") {
        std::cout << NetPimpl::TrainingNet::GetRequiredInputDimension() << ' '
                  << NetPimpl::RuntimeNet::GetRecommendedInputDimension(std::stoi(arg(2))) << std::endl;
        return 0;
    }
    if (cmd == "bench_infer") {
        // <anet> <H> <W> <tile> <reps>: annonet_infer() on a synthetic image, same timed region as annonet_infer_main.cpp:466-472
        const long h = std::stol(arg(3)), w = std::stol(arg(4));
        NetPimpl::RuntimeNet net;
        net.Deserialize(arg(2));
        NetPimpl::input_type img;
        synthetic_image(img, h, w, 1234);
        tiling::parameters tp;
        tp.max_tile_width = tp.max_tile_height = std::stoi(arg(5));
        tp.overlap_x = tp.overlap_y = NetPimpl::TrainingNet::GetRequiredInputDimension();
        annonet_infer_temp temp;
        dlib::matrix<uint16_t> result;
        const int reps = std::stoi(arg(6));
        std::vector<double> secs;
        for (int i = 0; i < reps; ++i) {
            const auto t0 = std::chrono::steady_clock::now();
            annonet_infer(net, img, result, temp, {}, {}, tp);
            secs.push_back(seconds_since(t0));
        }
        std::sort(secs.begin(), secs.end());
        const double med = secs[secs.size() / 2];
        std::cout << "{\"seconds\": " << med << ", \"mpix_per_s\": " << (h * w / med / 1e6) << ", \"threads\": " << host_threads()
                  << ", \"reps\": " << reps << "}" << std::endl;
        return 0;
    }
    if (cmd == "bench_train") {
        // <net> <N> <H> <W> <steps>: StartTraining x steps + sync on synthetic crops
        const long n = std::stol(arg(3)), h = std::stol(arg(4)), w = std::stol(arg(5));
        const int steps = std::stoi(arg(6));
        NetPimpl::TrainingNet t;
        t.Initialize();
        t.Deserialize(arg(2));
        t.SetLearningRate(0.1);
        std::vector<NetPimpl::input_type> imgs(n);
        std::vector<NetPimpl::training_label_type> labels(n);
        for (long i = 0; i < n; ++i) {
            synthetic_image(imgs[i], h, w, 99 + i);
            labels[i].set_size(h, w);
            for (long r = 0; r < h; ++r)
                for (long c = 0; c < w; ++c)
                    labels[i](r, c) = dlib::loss_multiclass_log_per_pixel_weighted_::weighted_label(static_cast<uint16_t>(((r / 16) + (c / 16)) % 2), 1.f);
        }
        const auto t0 = std::chrono::steady_clock::now();
        for (int s = 0; s < steps; ++s) t.StartTraining(imgs, labels);
        t.GetRuntimeNet();
        const double secs = seconds_since(t0);
        std::cout << "{\"seconds\": " << secs << ", \"crops_per_s\": " << (n * steps / secs) << ", \"threads\": " << host_threads()
                  << ", \"steps\": " << steps << "}" << std::endl;
        return 0;
    }
    return usage();
}
catch (const std::exception& e)
{
    std::cerr << "annonet_cli: " << e.what() << std::endl;
    return 1;
}
