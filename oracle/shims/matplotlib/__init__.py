"""TEST INFRASTRUCTURE ONLY -- empty stand-in so the reference's plotting
imports (sensor/noise_gen.py:6, simulation/rendering.py:1-2,
analysis/report.py:5-7) resolve in the build container. No plotting."""
