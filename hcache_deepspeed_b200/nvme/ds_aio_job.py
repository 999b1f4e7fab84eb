"""Run a benchmark command as a job, capturing its output (``deepspeed/nvme/ds_aio_job.py``)."""
import subprocess


class Job:

    def __init__(self, cmd_line, output_file=None, work_dir=None):
        self.cmd_line, self.output_file, self.work_dir = cmd_line, output_file, work_dir
        self.output_fd = None

    def cmd(self):
        return self.cmd_line

    def get_stdout(self):
        return self.output_fd

    def get_stderr(self):
        return self.output_fd

    def get_cwd(self):
        return self.work_dir

    def open_output_file(self):
        if self.output_file is not None:
            self.output_fd = open(self.output_file, "w")

    def close_output_file(self):
        if self.output_fd is not None:
            self.output_fd.close()
            self.output_fd = None


def run_job(job: Job) -> int:
    job.open_output_file()
    try:
        return subprocess.run(job.cmd(), cwd=job.get_cwd(), stdout=job.get_stdout(), stderr=job.get_stderr()).returncode
    finally:
        job.close_output_file()
