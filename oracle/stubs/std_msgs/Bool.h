#pragma once
namespace std_msgs { struct Bool { bool data; }; }
