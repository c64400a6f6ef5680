// scalix-b200: convenience alias kept for source compatibility
// (/root/reference scalix/filesystem.hpp:38-42).
#pragma once
#include <filesystem>
namespace sclx::filesystem {
using namespace std::filesystem;
}
