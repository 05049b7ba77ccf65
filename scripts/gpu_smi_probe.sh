#!/bin/bash
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap --format=csv,noheader,nounits -i 0 2>&1 | head -3
echo "---"
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_throttle_reasons.active,clocks_throttle_reasons.hw_slowdown,clocks_throttle_reasons.hw_thermal_slowdown,clocks_throttle_reasons.sw_thermal_slowdown,clocks_throttle_reasons.sw_power_cap --format=csv,noheader,nounits -i 0 2>&1 | head -3
echo "--- lms"
timeout 2 nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm --format=csv,noheader,nounits -i 0 -lms 200 2>&1 | head -5
