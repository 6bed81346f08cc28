"""Small Werkzeug-based stand-ins for `flask` / `flask_restful`, used only when those
packages are not installed (they are absent from the build image).  With Flask installed the
host shell imports the real thing."""
