// scalix-b200: exception helper with the reference's message format
// "<prefix><function> failed: <msg>" (/root/reference scalix/throw_exception.hpp:43-53).
#pragma once

#include <experimental/source_location>
#include <stdexcept>
#include <string>

namespace sclx {

template<class ExceptionType>
[[noreturn]] void throw_exception(
    const std::string& msg,
    const std::string& function_prefix = "sclx::",
    const std::experimental::source_location& location
    = std::experimental::source_location::current()
) {
    std::string text = function_prefix;
    text += location.function_name();
    text += " failed: ";
    text += msg;
    throw ExceptionType(text);
}

}  // namespace sclx
