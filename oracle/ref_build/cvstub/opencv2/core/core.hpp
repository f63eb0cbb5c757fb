// ORACLE reference build only (test infrastructure): adds the cv::FileStorage / cv::FileNode names that
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h mentions in its yaml save/load members on top of include/minicv.
// Those members are never called by the wrapper (the vocabulary is read with the reference's own
// loadFromTextFile); every stub aborts.
#ifndef ORACLE_CVSTUB_CORE_HPP
#define ORACLE_CVSTUB_CORE_HPP
#include_next <opencv2/core/core.hpp>
#include <cstdlib>
#include <sstream>
#include <iostream>
#include <string>
namespace cv {
struct FileNode {
    FileNode operator[](const char*) const { std::abort(); }
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](int) const { std::abort(); }
    size_t size() const { std::abort(); }
    operator int() const { std::abort(); }
    operator double() const { std::abort(); }
    operator std::string() const { std::abort(); }
};
struct FileStorage {
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char*, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string&) const { std::abort(); }
    void release() {}
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { std::abort(); return fs; }
}  // namespace cv
#endif
