// host_tool.cpp -- the C++ host-side readers and the option parser of include/nrc_host.hpp as a tiny CPU-only program, so that they can
// be tested without a GPU (tests/test_cpp_host.py):
//   host_tool hdr <file.hdr> <out.raw>        ReadFileHdr4f -> "<w> <h>" on stdout, float32 RGBA to <out.raw>
//   host_tool density <file> X Y Z <out.raw>  ReadFileDensity3D -> the floats written back
//   host_tool config [--key=value ...]        ParseNrcConfig -> one "key=value" line per field
//   host_tool pfm <out.pfm> | ppm <out.ppm>   a 3 x 2 test frame through WritePfm / WritePpm
#include <iostream>
#include "nrc_host.hpp"

int main(int argc, char** argv)
{
    try {
        const std::string cmd = argc > 1 ? argv[1] : "";
        if (cmd == "hdr" && argc == 4) {
            int w = 0, h = 0;
            std::vector<float> img = en::ReadFileHdr4f(argv[2], w, h);
            en::WriteFileDensity3D(argv[3], img);
            std::cout << w << " " << h << std::endl;
        } else if (cmd == "density" && argc == 7) {
            std::vector<float> d = en::ReadFileDensity3D(argv[2], (size_t)std::atol(argv[3]), (size_t)std::atol(argv[4]), (size_t)std::atol(argv[5]));
            en::WriteFileDensity3D(argv[6], d);
            std::cout << d.size() << std::endl;
        } else if (cmd == "config") {
            const en::NrcConfig c = en::ParseNrcConfig(argc - 1, argv + 1);
            std::cout << "width=" << c.width << "\nheight=" << c.height << "\ntrain=" << c.trainWidth << "x" << c.trainHeight << "\nframes=" << c.frames
                      << "\ndensity=" << c.density << "\ngrid=" << c.gridX << "," << c.gridY << "," << c.gridZ << "\nenvmap=" << c.envmap
                      << "\nuseNN=" << c.useNN << "\nspp=" << c.noNnSpp << "," << c.withNnSpp << "\ng=" << c.g << "\ndensityFactor=" << c.densityFactor
                      << "\ndir=" << c.dirZenith << "," << c.dirAzimuth << "," << c.dirStrength << "\npoint=" << c.pointPos[0] << "," << c.pointPos[1] << ","
                      << c.pointPos[2] << "," << c.pointStrength << "\nenv=" << c.envDirect << "," << c.envHpm << "\ncam=" << c.camPos[0] << "," << c.camPos[1]
                      << "," << c.camPos[2] << "\nfov=" << c.fovDeg << "\nlr=" << c.lr << "," << c.weightDecay << "," << c.beta1 << "\nmrhe=" << c.mrheLr << ","
                      << c.mrheWeightDecay << "," << c.levels << "," << c.hashTableSize << "," << c.minRes << "," << c.maxRes << "\nresolutions="
                      << c.resolutions << "\nprecision=" << c.precision << "\nfixed=" << c.fixedPreset << "\nadam=" << c.adam << "\ngpus=" << c.gpus
                      << "\nrank=" << c.rank << "\nout=" << c.out << "\noutPfm=" << c.outPfm << std::endl;
        } else if ((cmd == "pfm" || cmd == "ppm") && argc == 3) {
            const float rgba[24] = {0, 0.5f, 1, 1, 1, 2, 3, 1, 4, 5, 6, 1, 7, 8, 9, 1, 10, 11, 12, 1, 13, 14, 15, 1};
            const unsigned char bgra[24] = {3, 2, 1, 255, 6, 5, 4, 255, 9, 8, 7, 255, 12, 11, 10, 255, 15, 14, 13, 255, 18, 17, 16, 255};
            if (cmd == "pfm") en::WritePfm(argv[2], rgba, 3, 2);
            else en::WritePpm(argv[2], bgra, 3, 2);
        } else {
            std::cerr << "usage: host_tool hdr|density|config|pfm|ppm ..." << std::endl;
            return 2;
        }
    } catch (const std::exception& e) {
        std::cerr << "error: " << e.what() << std::endl;
        return 1;
    }
    return 0;
}
