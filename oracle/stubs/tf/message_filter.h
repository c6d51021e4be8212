#pragma once
// intentionally empty stand-in: the planner sources include this header but use nothing from it
