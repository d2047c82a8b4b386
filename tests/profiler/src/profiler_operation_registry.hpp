// Operation registry of ckProfiler: name -> (description, entry point); operations self-register from their own
// translation unit with REGISTER_PROFILER_OPERATION.  Role of profiler/src/profiler_operation_registry.hpp:12-79.
#pragma once
#include <functional>
#include <iostream>
#include <map>
#include <optional>
#include <string>

class ProfilerOperationRegistry final
{
    public:
    using Operation = std::function<int(int, char*[])>;
    static ProfilerOperationRegistry& GetInstance()
    {
        static ProfilerOperationRegistry registry;
        return registry;
    }
    std::optional<Operation> Get(const std::string& name) const
    {
        const auto it = entries_.find(name);
        if(it == entries_.end()) return std::nullopt;
        return it->second.second;
    }
    bool Add(const std::string& name, const std::string& description, Operation operation)
    {
        return entries_.emplace(name, std::make_pair(description, std::move(operation))).second;
    }
    friend std::ostream& operator<<(std::ostream& os, const ProfilerOperationRegistry& r)
    {
        os << "{\n";
        for(const auto& [name, entry] : r.entries_) os << "\t" << name << ": " << entry.first << "\n";
        return os << "}";
    }

    private:
    ProfilerOperationRegistry() = default;
    std::map<std::string, std::pair<std::string, Operation>> entries_;
};

#define CKP_CONCAT_IMPL(x, y) x##y
#define CKP_CONCAT(x, y) CKP_CONCAT_IMPL(x, y)
#define REGISTER_PROFILER_OPERATION(name, description, operation) \
    static const bool CKP_CONCAT(operation_registration_result_, __COUNTER__) = \
        ::ProfilerOperationRegistry::GetInstance().Add(name, description, operation)
